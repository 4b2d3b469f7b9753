// ComputeLevelSet (S3D.cpp:255-420) on sm_100a: per-cell seed from the cell-sorted particles, then the
// reference's 8 in-place Gauss-Seidel sweeps reproduced EXACTLY as block wavefronts.
//
// In the sweep with direction signs (sx,sy,sz) a cell only reads its three upwind neighbours, after
// they were finalised in the same sweep, and its three offers stay in x,y,z order inside one thread.
// Any schedule that respects that partial order gives the reference's result bit for bit, so the grid is
// cut into 16x16x8 tiles that run as a dataflow (see k_ls_sweep_dataflow) and inside a tile the 38 cell
// hyperplanes are walked by one 128-thread CTA with the tile + its upwind halo held in shared memory.
// Each cell carries its closest particle as (n*X, n*Y, n*Z, id): the candidate's position travels with
// the index, so the dependent chain contains no global gather, and phi is recomputed from it
// (ptDistance is symmetric in its arguments, S3D.cpp:234-236).
#include "fs3_internal.cuh"
#include <vector>
#include <algorithm>
#include <cstdio>
#include <cstdlib>


__device__ __forceinline__ float ls_dist(float cx, float cy, float cz, float4 q, float radius) {
	float dx = __fsub_rn(cx, q.x), dy = __fsub_rn(cy, q.y), dz = __fsub_rn(cz, q.z);
	float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
	return __fsub_rn(sqrtf(s), radius);
}

// ---- DIAMOND layout of the per-cell records -----------------------------------------------------------------------------
// A sweep step of one warp touches 32 cells on a diagonal of an (x, 32 rows) strip: x + r = const when the sweep runs the same
// way in x and y, x - r = const when it does not.  In natural order those are 32 different cache lines per load or store, and
// the load/store unit serialises them (measured: the whole sweep was bound by that).  A layout that makes one family of
// diagonals contiguous scatters the other, so the strip is cut into DIAMONDS instead: with u = x + r, v = x - r + 31, a 512-byte
// block holds the 32 cells of one 8 x 8 square in (u, v) (u + v is always odd, so half the square's points are cells).  Either
// kind of diagonal then crosses 8 blocks and uses 4 consecutive-in-time cells of each: 8 lines per warp instruction for every
// sweep direction, and every fetched sector is used.  u >> 3 and v >> 3 differ by at most 4, so a strip is (S + 7) / 8 block
// columns of 9 blocks, S = nx + 31 (1.27x the natural size at nx = 256).  Strips are ordered (z, y / 32): z planes stay contiguous.
struct LsLayout {
	int nty, ub;        // strips per plane, block columns per strip
	uint32_t strip;     // records per strip
};
static inline LsLayout ls_layout(int nx, int ny) {
	LsLayout L;
	L.nty = (ny + 31) / 32; L.ub = (nx + 31 + 7) / 8; L.strip = (uint32_t)L.ub * 9u * 32u;
	return L;
}
size_t fs3i_levelset_plane_elems(int nx, int ny) { const LsLayout L = ls_layout(nx, ny); return (size_t)L.nty * L.strip; }
size_t fs3i_levelset_rec_elems(int nx, int ny, int nz) {
	const size_t diamond = fs3i_levelset_plane_elems(nx, ny) * (size_t)nz, natural = (size_t)nx * ny * nz;
	return diamond > natural ? diamond : natural;
}
__device__ __forceinline__ uint32_t ls_off_uv(int u, int v) {   // block (U, V) at U * 9 + (V - U + 4) = 8 U + V + 4
	return (uint32_t)((((u & ~7) + (v >> 3) + 4) << 5) + ((((u & 7) << 3) + (v & 7)) >> 1));
}
__device__ __forceinline__ uint32_t ls_off(int x, int r) { return ls_off_uv(x + r, x - r + 31); }   // offset of cell (x, row r) inside its strip
__device__ __forceinline__ size_t ls_strip_base(const LsLayout& L, int ty, int z) { return ((size_t)z * L.nty + ty) * (size_t)L.strip; }

// Seed (S3D.cpp:281-300): minimum kernel distance over the particles of the cell (its 8 half-cells); the
// reference's strict `phi > k` in index order keeps the lowest index on ties.
// A cell's particles are four runs (FS3_FOR_CELL_RANGES) of about two particles each.  The eight run bounds and then the first
// two particles of every run are fetched before anything is compared (8 independent loads in flight per thread instead of
// one dependent load per loop iteration); longer runs finish in a tail loop.  min by (k, id) does not depend on the order.
__global__ void __launch_bounds__(256) k_ls_seed(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, float radius,
                                                 float4* __restrict__ rec, LsLayout L, size_t c0, size_t c1) {
	size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= c1) return;
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	float4 best = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	float bk = 0.f;
	int bid = -1;
	auto consider = [&](const float4 p) {
		float4 q = make_float4(__fmul_rn(p.x, (float)d.nx), __fmul_rn(p.y, (float)d.ny), __fmul_rn(p.z, (float)d.nz), p.w);
		float k = ls_dist((float)x, (float)y, (float)z, q, radius);
		int id = __float_as_int(p.w);
		// reference: particles in index order, replace on strictly smaller distance (S3D.cpp:296) == min by (k, id)
		if (bid < 0 || k < bk || (k == bk && id < bid)) { bk = k; bid = id; best = q; }
	};
	uint32_t B[4], E[4];
#pragma unroll
	for (int hr = 0; hr < 4; hr++) {
		const size_t row = (size_t)(2 * x) + (size_t)(2 * d.nx) * ((size_t)(2 * y + (hr & 1)) + (size_t)(2 * d.ny) * (2 * z + (hr >> 1)));
		B[hr] = __ldg(start + row); E[hr] = __ldg(start + row + 2);
	}
	float4 P[8];
#pragma unroll
	for (int hr = 0; hr < 4; hr++) {
		if (B[hr] < E[hr]) P[2 * hr] = __ldg(pos + B[hr]);
		if (B[hr] + 1 < E[hr]) P[2 * hr + 1] = __ldg(pos + B[hr] + 1);
	}
#pragma unroll
	for (int hr = 0; hr < 4; hr++) {
		if (B[hr] < E[hr]) consider(P[2 * hr]);
		if (B[hr] + 1 < E[hr]) consider(P[2 * hr + 1]);
	}
#pragma unroll
	for (int hr = 0; hr < 4; hr++)
		for (uint32_t t = B[hr] + 2; t < E[hr]; t++) consider(__ldg(pos + t));
	rec[ls_strip_base(L, y >> 5, z) + ls_off(x, y & 31)] = best;
}

// phi and closestParticles[] from the records the last sweep left.  Cells never reached keep their previous phi, as in the
// reference where m_Phi is only written together with closestParticles (S3D.cpp:248-251,296-299).
__global__ void __launch_bounds__(256) k_ls_finish(const float4* __restrict__ rec, LsLayout L, GridDims d, float radius, float* __restrict__ phi,
                                                   int* __restrict__ closest, size_t c0, size_t c1) {
	size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= c1) return;
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	const float4 q = __ldcg(rec + ls_strip_base(L, y >> 5, z) + ls_off(x, y & 31));
	const int id = __float_as_int(q.w);
	closest[c] = id;
	if (id >= 0) phi[c] = ls_dist((float)x, (float)y, (float)z, q, radius);
}

// ------------------------------------------------------------------------------------------------------------------------
// The 8 Gauss-Seidel sweeps (S3D.cpp:306-416), EXACTLY, as a pipeline of skewed warp strips.
//
// In the sweep with direction signs (dx,dy,dz) cell (a,b,c) -- sweep coordinates, 0 = most upwind -- reads the final values of
// (a-1,b,c), (a,b-1,c), (a,b,c-1) in that order; any schedule that respects this partial order reproduces the reference bit for
// bit.  The schedule here: a TILE is all of x times 32 rows (y) times BZ planes (z); BZ COMPUTE WARPS work on it in lock step,
// lane l of warp w handling cell  a = t - l - w  of row l, plane w  at step t.  Then
//   * the x candidate is the lane's own result of the previous step (a register),
//   * the y candidate is lane l-1's result of the previous step (one warp shuffle),
//   * the z candidate is warp w-1's result of the previous step (the shared-memory ring the results leave through),
// and a step's dependent chain holds no global memory access and no address arithmetic: the CTA is WARP-SPECIALISED.
//   * LW loader warps run ahead of the compute warps: they fetch the cells' records and the halo records a tile edge needs (row
//     -1 for lane 0, plane -1 for warp 0) with cp.async.cg into an LS_NS-stage shared-memory ring and hand every stage over
//     through a named barrier (loaders bar.arrive, compute warps bar.sync -- the same barrier that orders the compute warps' own
//     step-to-step hand-off, so a step costs ONE barrier);
//   * a storer warp drains the result ring to global memory (only records that changed), and publishes the tile's progress --
//     fence + release store of "steps stored" every LS_CH steps: the thread that fences is the thread that stored;
//   * a poller warp follows the progress words of the upwind tiles and posts them in shared memory for the loaders.
// (The first version did all of this in the compute warps: 770 ns per step, the serial latency of ~220 instructions of one
// warp; the same tiles and hand-offs, so the comparison is clean.)
// Tiles hand off to their downwind y / z neighbours at STEP granularity: the consumer runs the same skew and only has to stay
// 32 (y) or BZ (z) steps + the loaders' lead behind the producer's stored steps.  A whole sweep therefore costs about
// nx + (ny + nz) cells of dependent steps plus one L2 round trip per tile hop, instead of one complete-tile latency per hop.
// (sweep, tile) items are drawn by ticket in the order of their depth in the dependency DAG: every dependency of an item holds a
// lower ticket, so some running CTA can always proceed -- no co-residency assumption.
// Progress words: prog[tile] = (E0 + m) * 2048 + steps stored in sweep m, 2047 = sweep complete (wrap-safe comparisons).
// Conditions for (m, T): T and its four y/z neighbours have completed sweep m-1 (own cells final; nobody still reads what T is
// about to overwrite); the upwind y / z tiles are far enough into sweep m (checked by the loader of the halos).
// phi and closestParticles[] are written from the final records by k_ls_finish.
#define LS_CHK 4       // steps per loader chunk (one cp.async group, one hand-over barrier)
#define LS_NSLOT 4     // chunk slots of the input ring
#define LS_NS (LS_CHK * LS_NSLOT)   // stages of the input ring
#define LS_NO 8        // slots of the result ring
#define LS_LW 4        // loader warps
#define LS_CH 8        // steps between progress publications
#define LS_PROG_SHIFT 11
#define LS_PROG_DONE 2047u

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
	unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
	uint32_t v;
	asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
	uint32_t v;
	asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}
// Waits until *f has reached `need` (wrap-safe) and returns the value seen; polls with relaxed loads and a capped back-off, then
// takes the acquire.  ok = false: gave up (another CTA aborted, or ~2 s without progress: never hang the GPU).
__device__ __forceinline__ uint32_t ls_wait(const uint32_t* f, uint32_t need, int* flags, unsigned sleep_max, bool& ok) {
	uint32_t v = ld_acquire_u32(f);
	if ((int32_t)(v - need) >= 0) return v;
	long long t0 = clock64();
	unsigned ns = 32;
	for (unsigned it = 1;; it++) {
		v = ld_relaxed_u32(f);
		if ((int32_t)(v - need) >= 0) break;
		__nanosleep(ns);
		ns = min(ns * 2, sleep_max);
		if ((it & 255u) == 0) {
			if (*reinterpret_cast<volatile int*>(flags + 1)) { ok = false; return v; }
			if (clock64() - t0 > 4000000000LL) { flags[1] = 1; ok = false; return v; }
		}
	}
	return ld_acquire_u32(f);
}

__device__ __forceinline__ float ls_d2(float cx, float cy, float cz, float4 q) {
	float dx = __fsub_rn(cx, q.x), dy = __fsub_rn(cy, q.y), dz = __fsub_rn(cz, q.z);
	return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));   // ptDistance's radicand, S3D.cpp:234-236
}

#define LS_INF __int_as_float(0x7f800000)
#ifndef LS_EXP
#define LS_EXP 0    // timing experiments of the debug trace build only (wrong results): 1 = no exact fallback, 2 = helpers sleep while they wait, 4 = no arithmetic
#endif
#define LS_SPIN_PAUSE() do { if (LS_EXP & 2) __nanosleep(40); } while (0)
#ifdef FS3_LS_TRACE
__device__ unsigned long long* g_ls_trace;   // 4 timestamps per work item (debug side build only, tools/ls_trace.py)
__device__ __forceinline__ unsigned long long ls_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define LS_MARK(k) do { if (tid == 0) g_ls_trace[(size_t)tk * 8 + (k)] = ls_now(); } while (0)
#else
#define LS_MARK(k) do { } while (0)
#endif
#define LS_YH 32       // stages of the row -1 ring (16 bytes per compute warp and step: cheap to make deep)
// keeps the compiler from moving shared-memory accesses across a counter hand-over (the counters themselves are volatile)
#define LS_CFENCE() asm volatile("" ::: "memory")
template <int BZ> struct LsSmem {
	float4 own[LS_NS][BZ][32];   // the cells' records before the sweep, stage = step % LS_NS
	float4 zh[LS_NS][32];        // plane -1 of the tile (warp 0's z candidates)
	float4 yh[LS_YH][BZ];        // row -1 of the tile (lane 0's y candidates)
	float4 out[LS_NO][BZ][32];   // results, slot = step % LS_NO: read by the warp of the next plane one step later and by the storer
	uint32_t omask[LS_NO][BZ];   // lanes whose record changed (the storer writes only those)
	// hand-over counters (volatile accesses): steps completed per compute warp, chunks loaded per loader, steps written out by the
	// storer, steps the upwind tiles have stored in this sweep (posted by the poller), "a wait gave up: stop honouring waits"
	int cnt[BZ], loaded[LS_LW], stored, ky, kz, abort;
};
__device__ __forceinline__ int ld_volatile_s32(const int* p) { return *reinterpret_cast<const volatile int*>(p); }
__device__ __forceinline__ void st_volatile_s32(int* p, int v) { *reinterpret_cast<volatile int*>(p) = v; }

template <int BZ>
__global__ void __launch_bounds__((BZ + LS_LW + 2) * 32) k_ls_sweep_strips(float4* __restrict__ rec, LsLayout L, GridDims d, float radius, const uint32_t* __restrict__ items,
                                                                   uint32_t* prog, uint32_t* ticket, uint32_t nitems, int nty, int ntz, uint32_t E0,
                                                                   int* __restrict__ flags, unsigned sleep_max) {
	extern __shared__ float4 ls_smem_raw[];
	LsSmem<BZ>& sm = *reinterpret_cast<LsSmem<BZ>*>(ls_smem_raw);
	__shared__ uint32_t s_ticket;
	__shared__ int s_dead;
	int tid;
	asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tid));   // (volatile: keeps the compiler from re-reading the special register in the loops)
	const int l = tid & 31, w = tid >> 5;
	const float4 none = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	const bool fast_ok = radius == 1.0f;
	const int Ttot = d.nx + 31 + BZ - 1;      // steps of a tile: the last lane of the last warp handles a = nx-1 at step Ttot-1
	constexpr int NC = BZ * 32;
	static_assert(BZ <= 32 && LS_LW <= 32, "counters are gathered by one warp");
	// steps every compute warp has completed (one counter per lane, warp-wide minimum)
	auto steps_done = [&]() -> int { return __reduce_min_sync(0xffffffffu, l < BZ ? ld_volatile_s32(&sm.cnt[l]) : 0x7fffffff); };
	if (tid == 0) s_dead = 0;
	for (;;) {
		__syncthreads();   // the previous item's shared memory is no longer read
		if (tid == 0) { s_ticket = s_dead ? nitems : atomicAdd(ticket, 1u); sm.stored = 0; sm.abort = 0; }
		if (tid < BZ) sm.cnt[tid] = 0;
		if (tid < LS_LW) sm.loaded[tid] = 0;
		__syncthreads();
		const uint32_t tk = s_ticket;
		if (tk >= nitems) break;
		LS_MARK(0);
		const uint32_t it = __ldg(items + tk);
		const int m = (int)(it >> 27), Ty = (int)((it >> 9) & 511u), Tz = (int)((it >> 18) & 511u);
		const int dx = (m & 1) ? -1 : 1, dy = (m & 2) ? -1 : 1, dz = (m & 4) ? -1 : 1;   // sweep order of S3D.cpp:306-416
		const uint32_t base = (E0 + (uint32_t)m) << LS_PROG_SHIFT;
		const int y0 = Ty * 32, z0 = Tz * BZ;
		const int h = min(32, d.ny - y0), dep = min(BZ, d.nz - z0);
		uint32_t* const myprog = prog + ((size_t)Ty + (size_t)nty * Tz);
		const int uyT = Ty - dy, uzT = Tz - dz;                       // upwind tiles of this sweep
		const bool has_uy = uyT >= 0 && uyT < nty, has_uz = uzT >= 0 && uzT < ntz;
		const int gx0 = dx > 0 ? 0 : d.nx - 1;                       // physical x of a = 0
		// ---- this tile and its four neighbours have completed the previous sweep -------------------------------------------
		bool ok = true;
		if (tid < 5) {
			const int qy = Ty + (tid == 1 ? -1 : tid == 2 ? 1 : 0), qz = Tz + (tid == 3 ? -1 : tid == 4 ? 1 : 0);
			if (qy >= 0 && qy < nty && qz >= 0 && qz < ntz) (void)ls_wait(prog + ((size_t)qy + (size_t)nty * qz), base - 1u, flags, sleep_max, ok);
		}
		if (tid == 0) { sm.ky = has_uy ? 0 : 0x7fffffff; sm.kz = has_uz ? 0 : 0x7fffffff; }
		if (tid < NC && w < BZ - 1) sm.out[LS_NO - 1][w][l] = none;   // "results of step -1": what warp w+1 reads at step 0
		if (!__syncthreads_and(ok)) break;
		LS_MARK(1);
		if (w < BZ) {
			// ================================ compute warps ===============================================================
			// No CTA barrier in the step loop: warp w only needs warp w-1's result of the previous step, the loader of its rows and the
			// loader of the halos, and a free result slot -- four counters in shared memory, and the ring slack absorbs jitter.
			const bool lane_in = l < h && w < dep;
			const int gy = y0 + (dy > 0 ? l : h - 1 - l), gz = z0 + (dz > 0 ? w : dep - 1 - w);
			const float cyf = (float)gy, czf = (float)gz;
			const int* const cnt_up = &sm.cnt[w > 0 ? w - 1 : 0];
			const int* const cnt_down = &sm.cnt[w < BZ - 1 ? w + 1 : w];
			float4 prev = none;
			LS_MARK(2);
#pragma unroll 1
			for (int t = 0; t < Ttot; t++) {
				if ((t & (LS_CHK - 1)) == 0) {
					const int c = t / LS_CHK;   // the chunk's records and halos are in shared memory
					while (ld_volatile_s32(&sm.loaded[w & (LS_LW - 1)]) <= c || ld_volatile_s32(&sm.loaded[0]) <= c) { }
					LS_CFENCE();
				}
				// (read early, checked before the result is written: the latency hides behind the step)
				int stored = ld_volatile_s32(&sm.stored);
				int up = w > 0 ? ld_volatile_s32(cnt_up) : 0x7fffffff;
				int down = w < BZ - 1 ? ld_volatile_s32(cnt_down) : 0x7fffffff;
				const int slot = t & (LS_NS - 1);
				const int a = t - l - w;
				const bool act = lane_in && (unsigned)a < (unsigned)d.nx;
				const float cxf = (float)(gx0 + dx * a);
				const float4 mine0 = sm.own[slot][w][l];
				float4 Y;
				Y.x = __shfl_up_sync(0xffffffffu, prev.x, 1); Y.y = __shfl_up_sync(0xffffffffu, prev.y, 1);
				Y.z = __shfl_up_sync(0xffffffffu, prev.z, 1); Y.w = __shfl_up_sync(0xffffffffu, prev.w, 1);
				if (l == 0) Y = sm.yh[t & (LS_YH - 1)][w];
				const float4 X = prev;
				// clsInner on the upwind x-, y-, z-neighbour in that order (S3D.cpp:310-315 and siblings): a candidate with index <= 0
				// is never propagated (:242); a cell without a record takes any candidate and strict `dist < phi` keeps the
				// incumbent (:248) -- with "no record" and "no candidate" at distance +inf both are one comparison.
				const float m20 = __float_as_int(mine0.w) >= 0 ? ls_d2(cxf, cyf, czf, mine0) : LS_INF;
				const float x2 = __float_as_int(X.w) > 0 ? ls_d2(cxf, cyf, czf, X) : LS_INF;
				const float y2 = __float_as_int(Y.w) > 0 ? ls_d2(cxf, cyf, czf, Y) : LS_INF;
				// warp w-1 has finished step t-1: its result is this warp's z candidate
				while (up < t) up = ld_volatile_s32(cnt_up);
				LS_CFENCE();
				float4 Z;
				if (w > 0) Z = sm.out[(t + LS_NO - 1) & (LS_NO - 1)][w - 1][l]; else Z = sm.zh[slot][l];
				const float z2 = __float_as_int(Z.w) > 0 ? ls_d2(cxf, cyf, czf, Z) : LS_INF;
				float4 mine = mine0;
				float m2 = m20;
				bool changed = false, amb = false;
				// Fast path, branch-free, in squared distances (radius == 1 only): q2 >= m2 can never win (sqrt and "- radius" are
				// monotonic); q2 < m2 by more than 2^-21 relative with sqrt(m2) >= 0.5 always wins (the sqrt values differ by
				// >= 1 ulp and sqrt(m2) - 1 is exact); and below that -- both roots < 0.5, the common case inside the fluid --
				// q2 < m2 - 2^-22 always wins (the roots then differ by > 2^-22 - 2^-25, and rounding root - 1 to the 2^-24 grid of
				// (-1, -0.5] moves each side by at most 2^-25).  A comparison that falls between "never" and "always" marks the
				// step ambiguous and the whole warp repeats it with the reference's full expression (warp-uniform branch, rare).
#define LS_FAST(Q, q2)                                                                                                        \
	{                                                                                                                        \
		const bool take = q2 < (m2 >= 0.25f ? __fmul_rn(m2, 0.99999905f) : __fsub_rn(m2, 2.5e-7f)) && fast_ok;                \
		amb |= !take && q2 < m2;                                                                                             \
		mine.x = take ? Q.x : mine.x; mine.y = take ? Q.y : mine.y; mine.z = take ? Q.z : mine.z; mine.w = take ? Q.w : mine.w; \
		m2 = take ? q2 : m2; changed |= take;                                                                                \
	}
				if (!(LS_EXP & 4)) {
				LS_FAST(X, x2)
				LS_FAST(Y, y2)
				LS_FAST(Z, z2)
				}
#undef LS_FAST
				if (!(LS_EXP & 1) && __any_sync(0xffffffffu, amb && act)) {
					mine = mine0; m2 = m20; changed = false;
#define LS_TRY(Q, q2)                                                                                                         \
	{                                                                                                                        \
		const bool take = __fsub_rn(sqrtf(q2), radius) < __fsub_rn(sqrtf(m2), radius);                                       \
		mine.x = take ? Q.x : mine.x; mine.y = take ? Q.y : mine.y; mine.z = take ? Q.z : mine.z; mine.w = take ? Q.w : mine.w; \
		m2 = take ? q2 : m2; changed |= take;                                                                                \
	}
					LS_TRY(X, x2)
					LS_TRY(Y, y2)
					LS_TRY(Z, z2)
#undef LS_TRY
				}
				if (!act) mine = none;
				const int oslot = t & (LS_NO - 1);
				// the result slot is free: the storer (which walks the warps' skew, iteration i = step + warp) has written out this
				// warp's step t - LS_NO, and warp w+1 has read it (in its step t - LS_NO + 1)
				while ((stored < t - LS_NO + 1 + w || down < t - LS_NO + 2) && !ld_volatile_s32(&sm.abort)) {
					stored = ld_volatile_s32(&sm.stored);
					if (w < BZ - 1) down = ld_volatile_s32(cnt_down);
				}
				LS_CFENCE();
				sm.out[oslot][w][l] = mine;
				const unsigned cm = __ballot_sync(0xffffffffu, act && changed);
				if (l == 0) sm.omask[oslot][w] = cm;
				prev = mine;
				LS_CFENCE();
				__syncwarp();
				if (l == 0) st_volatile_s32(&sm.cnt[w], t + 1);
			}
			LS_MARK(3);
		} else if (w < BZ + LS_LW) {
			// ================================ loader warps =================================================================
			// loader j fetches the records of compute warps j, j + LS_LW, ...; loader 0 also row -1 (lanes 0..BZ-1, one per compute
			// warp) and plane -1 (all lanes), which depend on the upwind tiles' progress
			const int j = w - BZ;
			const bool row_in = l < h;
			const int r = (y0 + (dy > 0 ? l : h - 1 - l)) & 31;
			const int gzh = dz > 0 ? z0 - 1 : z0 + dep, rY = dy > 0 ? 31 : 0;
			const float4* const stripZ = rec + ls_strip_base(L, Ty, has_uz ? gzh : z0);
			// lane k < BZ of loader 0: row -1 of plane k
			const int gzk = z0 + (dz > 0 ? l : dep - 1 - l);
			const float4* const stripY = rec + ls_strip_base(L, has_uy ? uyT : Ty, (l < dep) ? gzk : z0);
			int seen = 0;   // steps every compute warp has completed, as last read
			const int nch = (Ttot + LS_CHK - 1) / LS_CHK;
			// u, v of the diamond layout for column a = 0 of this lane's row; a column step moves both by dx
			const int u0 = gx0 + r, v0 = gx0 - r + 31;
#pragma unroll 1
			for (int c = 0; c < nch + 2; c++) {
				if (c < nch) {
					// the chunk slot was last used by chunk c - LS_NSLOT: the warps this loader serves have completed its steps (loader 0:
					// that includes warp 0, the reader of plane -1; the deep row -1 ring only needs EVERY warp within LS_YH steps)
					while (seen < LS_CHK * (c - LS_NSLOT + 1)) {
						LS_SPIN_PAUSE();
						seen = 0x7fffffff;
#pragma unroll
						for (int cw = j; cw < BZ; cw += LS_LW) seen = min(seen, ld_volatile_s32(&sm.cnt[cw]));
					}
					if (j == 0) { int all = steps_done(); while (all < LS_CHK * (c + 1) - LS_YH) all = steps_done(); }
					LS_CFENCE();
					const int tp0 = c * LS_CHK;
					if (j == 0) {
						// the upwind tiles have STORED the steps that produce what the chunk's last step reads: row 31 / plane BZ-1 of
						// the same column, 32 / BZ steps later in the producer's own skew
						const int needy = min(tp0 + LS_CHK - 1 + 32, Ttot), needz = min(tp0 + LS_CHK - 1 + BZ, Ttot);
						while ((ld_volatile_s32(&sm.ky) < needy || ld_volatile_s32(&sm.kz) < needz) && !ld_volatile_s32(&sm.abort)) { LS_SPIN_PAUSE(); }
#pragma unroll
						for (int k = 0; k < LS_CHK; k++) {
							const int tp = tp0 + k, slot = tp & (LS_NS - 1);
							const int az = tp - l;
							if (has_uz && row_in && (unsigned)az < (unsigned)d.nx) cp_async16(&sm.zh[slot][l], stripZ + ls_off_uv(u0 + dx * az, v0 + dx * az)); else sm.zh[slot][l] = none;
							if (l < BZ) {
								const int ay = tp - l;   // compute warp l handles column tp - l in its lane 0
								if (has_uy && l < dep && (unsigned)ay < (unsigned)d.nx) cp_async16(&sm.yh[tp & (LS_YH - 1)][l], stripY + ls_off(gx0 + dx * ay, rY)); else sm.yh[tp & (LS_YH - 1)][l] = none;
							}
						}
					}
#pragma unroll
					for (int cw = j; cw < BZ; cw += LS_LW) {
						const int gz = z0 + (dz > 0 ? cw : dep - 1 - cw);
						const float4* const strip = rec + ls_strip_base(L, Ty, cw < dep ? gz : z0);
						const bool pl_in = row_in && cw < dep;
#pragma unroll
						for (int k = 0; k < LS_CHK; k++) {
							const int tp = tp0 + k, slot = tp & (LS_NS - 1);
							const int a = tp - l - cw;
							if (pl_in && (unsigned)a < (unsigned)d.nx) cp_async16(&sm.own[slot][cw][l], strip + ls_off_uv(u0 + dx * a, v0 + dx * a));
							else sm.own[slot][cw][l] = none;
						}
					}
				}
				cp_async_commit();
				cp_async_wait<2>();
				// chunk c - 2 (records and none fills of every lane) is in shared memory
				LS_CFENCE();
				__syncwarp();
				if (l == 0 && c >= 2) st_volatile_s32(&sm.loaded[j], c - 1);
			}
		} else if (w == BZ + LS_LW) {
			// ================================ storer warp ==================================================================
			// lane l writes the changed records of row l for every plane, and the warp publishes "steps stored"
			const int r = (y0 + (dy > 0 ? l : h - 1 - l)) & 31;
			const int u0 = gx0 + r, v0 = gx0 - r + 31;
			// The storer walks the compute warps' skew: iteration i writes step i - cw of warp cw, which that warp completes about
			// when warp 0 completes step i -- so no warp is held back by the ring for the sake of a slower one.
			int pub = 0;
#pragma unroll 1
			for (int i = 0; i < Ttot + BZ - 1; i++) {
				// warp cw has finished step i - cw (lanes 0..BZ-1 look at one counter each)
				for (;;) {
					const int c = l < BZ ? ld_volatile_s32(&sm.cnt[l]) : 0x7fffffff;
					if (__all_sync(0xffffffffu, l >= BZ || c > min(i - l, Ttot - 1))) break;
					LS_SPIN_PAUSE();
				}
				LS_CFENCE();
#pragma unroll
				for (int cw = 0; cw < BZ; cw++) {
					const int ts = i - cw;
					if ((unsigned)ts < (unsigned)Ttot) {
						const unsigned cm = sm.omask[ts & (LS_NO - 1)][cw];
						if ((cm >> l) & 1u) {
							const int a = ts - l - cw;
							const int gz = z0 + (dz > 0 ? cw : dep - 1 - cw);
							__stcg(rec + ls_strip_base(L, Ty, gz) + ls_off_uv(u0 + dx * a, v0 + dx * a), sm.out[ts & (LS_NO - 1)][cw][l]);
						}
					}
				}
				LS_CFENCE();
				__syncwarp();
				if (l == 0) {
					st_volatile_s32(&sm.stored, i + 1);
					const int done = i + 1 - (BZ - 1);   // steps of which EVERY warp's records are written out
					if (done >= pub + LS_CH && done < Ttot) { __threadfence(); st_release_u32(myprog, base + (uint32_t)done); pub = done; }
				}
			}
			__syncwarp();
			if (l == 0) { __threadfence(); st_release_u32(myprog, base + LS_PROG_DONE); }   // sweep m of this tile is complete
		} else {
			// ================================ poller warp ==================================================================
			if (l == 0) {
				const uint32_t* const prog_uy = prog + ((size_t)(has_uy ? uyT : Ty) + (size_t)nty * Tz);
				const uint32_t* const prog_uz = prog + ((size_t)Ty + (size_t)nty * (has_uz ? uzT : Tz));
				int ky = has_uy ? 0 : Ttot, kz = has_uz ? 0 : Ttot;
				long long t_last = clock64();
				// RELAXED polls: an acquire load is LDG + CCTL.IVALL on this architecture (an L1 invalidation per poll), and none
				// is needed -- the records these words guard are fetched with cp.async.cg, which bypasses L1, after they were posted
				while (ky < Ttot || kz < Ttot) {
					bool moved = false;
					if (ky < Ttot) {
						const int k = min((int)(ld_relaxed_u32(prog_uy) - base), Ttot);
						if (k > ky) { ky = k; st_volatile_s32(&sm.ky, k); moved = true; }
					}
					if (kz < Ttot) {
						const int k = min((int)(ld_relaxed_u32(prog_uz) - base), Ttot);
						if (k > kz) { kz = k; st_volatile_s32(&sm.kz, k); moved = true; }
					}
					if (moved) t_last = clock64();
					else {
						__nanosleep(200);
						if (clock64() - t_last > 4000000000LL || *reinterpret_cast<volatile int*>(flags + 1)) {
							flags[1] = 1; st_volatile_s32(&s_dead, 1); st_volatile_s32(&sm.abort, 1);   // waits are no longer honoured: the item runs out
							break;
						}
					}
				}
			}
		}
	}
}

// seed of the cells of this rank's slab (world == 1: all cells)
int fs3i_levelset_seed(fs3_sim* s, cudaStream_t st) {
	GridDims d{s->nx, s->ny, s->nz};
	if (!st) st = s->stream;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SEED], st);
	const size_t plane = (size_t)s->nx * s->ny, c0 = plane * s->z0, c1 = plane * s->z1;
	k_ls_seed<<<cdiv((long long)(c1 - c0), 256), 256, 0, st>>>(s->pos[s->cur], s->cell_start, d, s->prm.particle_radius, s->ls_rec, ls_layout(s->nx, s->ny), c0, c1);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

int fs3i_levelset(fs3_sim* s) {
	if (s->prm.levelset_mode == 1) return fs3i_levelset_lines(s);
	int rc = fs3i_bin(s);
	if (rc) return rc;
	if ((rc = fs3i_levelset_seed(s))) return rc;
	s->ls_overlapped = false;
	return fs3i_levelset_sweeps(s);
}

// the 8 sweeps over the WHOLE grid (a sharded handle runs them replicated on the gathered seed records)
#ifndef LS_BZ
#define LS_BZ 8     // planes per tile = compute warps per CTA
#endif
int fs3i_levelset_sweeps(fs3_sim* s, cudaStream_t st) {
	GridDims d{s->nx, s->ny, s->nz};
	float radius = s->prm.particle_radius;
	if (!st) st = s->stream;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SWEEP], st);
	const int nty = cdiv(d.ny, 32), ntz = cdiv(d.nz, LS_BZ);
	FS3_CUDA(cudaMemsetAsync(s->ls_ticket, 0, sizeof(uint32_t), st));
	const LsLayout L = ls_layout(s->nx, s->ny);
#ifdef FS3_LS_TRACE
	static unsigned long long* trace_dev = nullptr;
	const size_t trace_n = (size_t)8 * s->ls_ntiles * 8;
	if (!trace_dev) { FS3_CUDA(cudaMalloc((void**)&trace_dev, trace_n * 8)); FS3_CUDA(cudaMemcpyToSymbol(g_ls_trace, &trace_dev, sizeof(trace_dev))); }
	FS3_CUDA(cudaMemsetAsync(trace_dev, 0, trace_n * 8, st));
#endif
	k_ls_sweep_strips<LS_BZ><<<s->ls_blocks, (LS_BZ + LS_LW + 2) * 32, sizeof(LsSmem<LS_BZ>), st>>>(
	    s->ls_rec, L, d, radius, s->ls_order, s->ls_done, s->ls_ticket, 8u * (uint32_t)s->ls_ntiles, nty, ntz, s->ls_epoch, s->dev_flags,
	    (unsigned)s->ls_sleep_ns);
	FS3_LAUNCH_CHECK();
	k_ls_finish<<<cdiv((long long)s->nC, 256), 256, 0, st>>>(s->ls_rec, L, d, radius, s->phi, s->closest, 0, s->nC);
	FS3_LAUNCH_CHECK();
#ifdef FS3_LS_TRACE
	{
		std::vector<unsigned long long> h(trace_n);
		std::vector<uint32_t> ord((size_t)8 * s->ls_ntiles);
		FS3_CUDA(cudaStreamSynchronize(st));
		FS3_CUDA(cudaMemcpy(h.data(), trace_dev, trace_n * 8, cudaMemcpyDeviceToHost));
		FS3_CUDA(cudaMemcpy(ord.data(), s->ls_order, ord.size() * 4, cudaMemcpyDeviceToHost));
		if (FILE* fp = fopen("/tmp/ls_trace.bin", "wb")) { fwrite(h.data(), 8, trace_n, fp); fwrite(ord.data(), 4, ord.size(), fp); fclose(fp); }
	}
#endif
	s->ls_epoch += 8;
	s->closest_valid = true;
	return FS3_OK;
}

// The sweeps only touch ls_rec / phi / closest, which nothing else reads before the pressure system is set up: inside a step
// they overlap the P2G gather and the extrapolation (both bandwidth / issue bound, while the sweeps wait on tile dependencies).
int fs3i_levelset_fork(fs3_sim* s, bool after_barrier, bool with_seed) {
	FS3_CUDA(cudaEventRecord(s->ev_fork, s->stream));
	FS3_CUDA(cudaStreamWaitEvent(s->stream2, s->ev_fork, 0));
	int rc;
	s->ls_seed_forked = with_seed;
	if (with_seed && (rc = fs3i_levelset_seed(s, s->stream2))) return rc;   // single GPU: the seed needs the sorted particles only
	// sharded: the gathered seed records are complete once every rank has published the current barrier epoch; that wait
	// sits on stream2 too, so the main stream goes on with the P2G gather meanwhile
	if (after_barrier && (rc = fs3i_wait(s, s->stream2))) return rc;
	if ((rc = fs3i_levelset_sweeps(s, s->stream2))) return rc;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev_ls_end, s->stream2);
	FS3_CUDA(cudaEventRecord(s->ev_join, s->stream2));
	s->ls_forked = true;
	s->ls_overlapped = true;
	return FS3_OK;
}
int fs3i_levelset_join(fs3_sim* s) {
	if (!s->ls_forked) return FS3_OK;
	FS3_CUDA(cudaStreamWaitEvent(s->stream, s->ev_join, 0));
	s->ls_forked = false;
	return FS3_OK;
}

// Work-item table, progress words and the persistent grid size (called once from fs3_create).
// Items (sweep m, tile) are sorted by their level in the dependency DAG -- the earliest "time" at which the item could start
// if every hand-off took one unit -- which is a topological order (every dependency has a strictly smaller level), so ticket
// order is deadlock-free, and tiles of one wavefront hold neighbouring tickets, i.e. run at the same time.
__global__ void k_fill_u32(uint32_t* p, size_t n, uint32_t v) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < n) p[q] = v;
}
int fs3i_levelset_init(fs3_sim* s) {
	const int nty = cdiv(s->ny, 32), ntz = cdiv(s->nz, LS_BZ);
	if (nty > 511 || ntz > 511 || s->nx + 31 + LS_BZ >= (int)LS_PROG_DONE) { s->err = "grid too large for the level-set tile table"; return FS3_ERR_INVALID; }
	const int nt = nty * ntz;
	auto tid = [&](int y, int z) { return y + nty * z; };
	std::vector<int> prev(nt, -1), cur(nt, 0);
	std::vector<std::pair<int, uint32_t>> items;
	items.reserve((size_t)8 * nt);
	int maxlevel = 0;
	for (int m = 0; m < 8; m++) {
		const int dy = (m & 2) ? -1 : 1, dz = (m & 4) ? -1 : 1;
		for (int kz = 0; kz < ntz; kz++)
			for (int ky = 0; ky < nty; ky++) {   // sweep-space order: upwind neighbours are visited first
				const int y = dy > 0 ? ky : nty - 1 - ky, z = dz > 0 ? kz : ntz - 1 - kz;
				int L = 0;
				if (ky > 0) L = std::max(L, cur[tid(y - dy, z)] + 1);
				if (kz > 0) L = std::max(L, cur[tid(y, z - dz)] + 1);
				if (m > 0) {
					L = std::max(L, prev[tid(y, z)] + 1);
					if (y > 0) L = std::max(L, prev[tid(y - 1, z)] + 1);
					if (y < nty - 1) L = std::max(L, prev[tid(y + 1, z)] + 1);
					if (z > 0) L = std::max(L, prev[tid(y, z - 1)] + 1);
					if (z < ntz - 1) L = std::max(L, prev[tid(y, z + 1)] + 1);
				}
				cur[tid(y, z)] = L;
				maxlevel = std::max(maxlevel, L);
				items.emplace_back(L, ((uint32_t)y << 9) | ((uint32_t)z << 18) | ((uint32_t)m << 27));
			}
		prev = cur;
	}
	std::stable_sort(items.begin(), items.end(), [](const std::pair<int, uint32_t>& a, const std::pair<int, uint32_t>& b) { return a.first < b.first; });
	std::vector<uint32_t> ord(items.size());
	for (size_t i = 0; i < items.size(); i++) ord[i] = items[i].second;
	int rc;
	if ((rc = fs3i_alloc(s, (void**)&s->ls_order, sizeof(uint32_t) * ord.size()))) return rc;
	if ((rc = fs3i_alloc(s, (void**)&s->ls_done, sizeof(uint32_t) * nt))) return rc;
	if ((rc = fs3i_alloc(s, (void**)&s->ls_ticket, sizeof(uint32_t)))) return rc;
	FS3_CUDA(cudaMemcpyAsync(s->ls_order, ord.data(), sizeof(uint32_t) * ord.size(), cudaMemcpyHostToDevice, s->stream));
	// the first call's sweeps have global indices 8..15: every tile starts as "sweep 7 complete"
	s->ls_epoch = 8;
	k_fill_u32<<<cdiv(nt, 256), 256, 0, s->stream>>>(s->ls_done, (size_t)nt, (7u << LS_PROG_SHIFT) + LS_PROG_DONE); FS3_LAUNCH_CHECK();
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	s->ls_ntiles = nt;
	s->ls_levels = maxlevel + 1;
	int per_sm = 0, sms = 0;
	const size_t smem = sizeof(LsSmem<LS_BZ>);
	FS3_CUDA(cudaFuncSetAttribute(k_ls_sweep_strips<LS_BZ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	FS3_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ls_sweep_strips<LS_BZ>, (LS_BZ + LS_LW + 2) * 32, smem));
	FS3_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
	// Every tile of a sweep can be in flight at once (they follow each other at hand-off distance), so the grid holds up to one
	// CTA per tile; the ticket scheme needs no co-residency, any count is valid.
	s->ls_sleep_ns = 256;
	const long long resident = (long long)std::max(1, std::min(per_sm, 2)) * sms;
	s->ls_blocks = (int)std::max<long long>(1, std::min<long long>(resident, nt));
	return FS3_OK;
}
