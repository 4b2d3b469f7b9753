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

// Seed (S3D.cpp:281-300): minimum kernel distance over the particles of the cell (its 8 half-cells); the
// reference's strict `phi > k` in index order keeps the lowest index on ties.
// A cell's particles are four runs (FS3_FOR_CELL_RANGES) of about two particles each.  The eight run bounds and then the first
// two particles of every run are fetched before anything is compared (8 independent loads in flight per thread instead of
// one dependent load per loop iteration); longer runs finish in a tail loop.  min by (k, id) does not depend on the order.
__global__ void __launch_bounds__(256) k_ls_seed(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, float radius,
                                                 float4* __restrict__ rec, size_t c0, size_t c1) {
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
	rec[c] = best;
}

// All 8 sweeps in ONE persistent launch, scheduled as a dataflow over 16x16x8 tiles, ONE 128-THREAD CTA PER TILE.
// Tiles are a fixed partition of physical space.  A work item is (sweep m, tile); the host orders all
// 8*ntiles items by their earliest possible start level in the dependency DAG (fs3i_levelset_init), and
// CTAs draw items through a ticket counter.  A CTA that holds an item waits (polling per-tile
// completion epochs in L2) until
//   * the tile itself and its 6 face neighbours have completed sweep m-1 (RAW on its own cells, WAR on
//     the halo cells its neighbours read in sweep m-1), and
//   * its 3 upwind neighbours have completed sweep m (RAW on the halo),
// all of which hold lower tickets, so some running CTA can always make progress (no co-residency
// assumption, no grid-wide barrier), and consecutive sweeps pipeline behind each other as far as the
// DAG allows.  The sweep is bound by its critical path: 8 n/tx + 4 n/ty + 2 n/tz tile hand-offs (x flips
// between every pair of sweeps, y every second, z once), each costing a few L2 round trips (poll, halo
// load, store, release) plus the tile's own tx+ty+tz-2 cell hyperplanes at one shared-memory round trip
// each.  16x16x8 halves the number of hand-offs of a warp-sized 8x8x4 tile (256 vs 512 at 256^3) while a
// hyperplane step still costs one barrier.  The tile and its three upwind halo faces sit in shared memory
// in SWEEP space (index 0 = most upwind) so that one code path serves all 8 directions; 128 threads = the
// 16x8 x-lines of the tile, thread (b,c) marching along a with hyperplane index s = a+b+c.  Candidates are
// first compared by squared distance (sqrt and "- radius" are monotonic, so a candidate that is not
// closer in d^2 can never win the reference's `dist < phi` test); the full reference expression decides
// the rest.  Records move with cp.async.cg / st.cg (L2 only: L1 is not coherent across SMs).  The last
// sweep also writes phi and closestParticles[].
#define LS_TX 16
#define LS_TY 16
#define LS_TZ 8
#define LS_PX (LS_TX + 2)   // x pitch 18 (even): on a hyperplane a+b+c=s the 8 lanes of a quarter warp (consecutive b) hit 8 different 16-byte bank groups
#define LS_PY (LS_TY + 1)
#define LS_PZ (LS_TZ + 1)
#define LS_STEPS (LS_TX + LS_TY + LS_TZ - 2)
#define LS_LINES (LS_TY * LS_TZ)   // threads that march
#define LS_THREADS 256             // all of them load, run the quiet-tile test and write phi
#define LS_CELLS (LS_PX * LS_PY * LS_PZ)
#define LS_SMEM_BYTES (LS_CELLS * (sizeof(float4) + sizeof(float)))

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
	unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }
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
// Polls with relaxed loads and a capped exponential back-off, then takes the acquire once the epoch is there.
// false = gave up (another CTA aborted, or ~2 s without progress: never hang the GPU)
// sys: the word is written by another GPU (a neighbouring rank's boundary tile)
__device__ __forceinline__ bool ls_wait(const uint32_t* f, uint32_t need, int* flags, unsigned sleep_max, bool sys = false) {
	if ((int32_t)((sys ? fs3_ld_acquire_sys(f) : ld_acquire_u32(f)) - need) >= 0) return true;
	long long t0 = clock64();
	unsigned ns = 32;
	for (unsigned it = 1;; it++) {
		if ((int32_t)((sys ? fs3_ld_relaxed_sys(f) : ld_relaxed_u32(f)) - need) >= 0) break;
		__nanosleep(ns);
		ns = min(ns * 2, sleep_max);
		if ((it & 255u) == 0) {
			if (*reinterpret_cast<volatile int*>(flags + 1)) return false;
			if (clock64() - t0 > 4000000000LL) { flags[1] = 1; return false; }
		}
	}
	return (int32_t)((sys ? fs3_ld_acquire_sys(f) : ld_acquire_u32(f)) - need) >= 0;
}

// Where the sweeps' arrays live.  A z-slab sharded simulation sweeps the SAME global dataflow: tile layers are aligned to the
// slabs (layer Tz belongs to rank Tz / lps), every rank runs the items of its own tiles on its own copy of the arrays (global
// indexing), and a tile on a slab boundary hands its boundary plane of records and its completion epoch to the neighbouring
// rank by storing them into that rank's arrays (NVLink peer stores, fence, release) -- the consumer only ever polls and reads
// its own memory.  base[q] + off_x is array x of rank q (one arena layout for all ranks); a single-GPU handle has world = 1,
// base[0] = nullptr and absolute addresses as offsets.  A lockstep group (all ranks in one process on one device) runs the
// items of ALL ranks in one launch with the same code.
struct LsCtx {
	char* base[FS3_MAX_WORLD];
	size_t off_rec, off_done, off_phi, off_closest;
	int world, slab, lps;   // planes per slab, tile layers per slab
	int sys;                // 1: neighbouring ranks are other GPUs (system-scope flag accesses)
};

__device__ __forceinline__ float ls_d2(float cx, float cy, float cz, float4 q) {
	float dx = __fsub_rn(cx, q.x), dy = __fsub_rn(cy, q.y), dz = __fsub_rn(cz, q.z);
	return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));   // ptDistance's radicand, S3D.cpp:234-236
}

#ifdef FS3_LS_TRACE
__device__ unsigned long long* g_ls_trace;   // 8 timestamps per work item (debug builds only, tools/ls_trace.py)
__device__ __forceinline__ unsigned long long ls_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define LS_MARK(k) do { if (threadIdx.x == 0) g_ls_trace[(size_t)t * 8 + (k)] = ls_now(); } while (0)
#else
#define LS_MARK(k) do { } while (0)
#endif

#define LS_INF __int_as_float(0x7f800000)
__global__ void __launch_bounds__(LS_THREADS, 3) k_ls_sweep_dataflow(LsCtx C, GridDims d, float radius,
                                                                   const uint32_t* __restrict__ items, uint32_t* ticket,
                                                                   uint32_t nitems, int ntx, int nty, int ntz, uint32_t epoch0,
                                                                   int* __restrict__ flags, unsigned sleep_max) {
	extern __shared__ float4 ls_smem[];
	__shared__ uint32_t s_ticket;
	__shared__ unsigned s_planes[2];
	float4* const sm = ls_smem;
	float* const m2s = reinterpret_cast<float*>(ls_smem + LS_CELLS);   // squared distance of the cell to its current record
	const int tid = threadIdx.x;
#define SMI(a, b, c) (((a) + 1) + LS_PX * (((b) + 1) + LS_PY * ((c) + 1)))
	const float4 none = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	const size_t sY = (size_t)d.nx, sZ = (size_t)d.nx * d.ny;
	for (;;) {
		__syncthreads();   // the previous item's shared memory is no longer read
		if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
		__syncthreads();
		const uint32_t t = s_ticket;
		if (t >= nitems) break;
		LS_MARK(0);
		const uint32_t it = __ldg(items + t);
		const int m = (int)(it >> 27), Tx = (int)(it & 511u), Ty = (int)((it >> 9) & 511u), Tz = (int)((it >> 18) & 511u);
		const int dx = (m & 1) ? -1 : 1, dy = (m & 2) ? -1 : 1, dz = (m & 4) ? -1 : 1;   // sweep order of S3D.cpp:306-416
		const uint32_t E = epoch0 + (uint32_t)m + 1u;
		// tile layer Tz = layer `lay` of rank q's slab; the arrays are rank q's copies
		const int q = Tz / C.lps, lay = Tz - q * C.lps;
		float4* const rec = reinterpret_cast<float4*>(C.base[q] + C.off_rec);
		uint32_t* const done = reinterpret_cast<uint32_t*>(C.base[q] + C.off_done);
		const int x0 = Tx * LS_TX, y0 = Ty * LS_TY, z0 = q * C.slab + lay * LS_TZ;
		const int wx = min(LS_TX, d.nx - x0), wy = min(LS_TY, d.ny - y0), wz = min(LS_TZ, (q + 1) * C.slab - z0);
		const size_t tile_idx = (size_t)Tx + (size_t)ntx * ((size_t)Ty + (size_t)nty * Tz);
		uint32_t* const mydone = done + tile_idx;
		const bool peer_dn = lay == 0 && q > 0, peer_up = lay == C.lps - 1 && q < C.world - 1;   // neighbouring ranks below / above this tile
		// ---- own cells are final once this tile has completed sweep m-1: start their copy first ----------
		bool ok = true;
		if (tid == 0) ok = ls_wait(mydone, E - 1u, flags, sleep_max);
		if (!__syncthreads_and(ok)) break;
		LS_MARK(1);
#pragma unroll
		for (int i = 0; i < LS_TX * LS_TY * LS_TZ / LS_THREADS; i++) {
			const int e = tid + LS_THREADS * i, px = e % LS_TX, py = (e / LS_TX) % LS_TY, pz = e / (LS_TX * LS_TY);
			const int a = (px < wx && dx < 0) ? wx - 1 - px : px, b = (py < wy && dy < 0) ? wy - 1 - py : py, c = (pz < wz && dz < 0) ? wz - 1 - pz : pz;
			if (px < wx && py < wy && pz < wz) cp_async16(sm + SMI(a, b, c), rec + ((size_t)(x0 + px) + sY * (size_t)(y0 + py) + sZ * (size_t)(z0 + pz)));
			else sm[SMI(a, b, c)] = none;
		}
		// ---- the six neighbours: upwind ones must have finished THIS sweep, the others the previous one ---
		if (tid >= 1 && tid <= 6) {
			const int ax = (tid - 1) >> 1, sgn = (tid & 1) ? -1 : 1;   // threads 1,2: x-,x+   3,4: y-,y+   5,6: z-,z+
			const int qx = Tx + (ax == 0 ? sgn : 0), qy = Ty + (ax == 1 ? sgn : 0), qz = Tz + (ax == 2 ? sgn : 0);
			const int dir = ax == 0 ? dx : (ax == 1 ? dy : dz);
			if (qx >= 0 && qx < ntx && qy >= 0 && qy < nty && qz >= 0 && qz < ntz)
				ok = ls_wait(done + ((size_t)qx + (size_t)ntx * ((size_t)qy + (size_t)nty * qz)), sgn == -dir ? E : E - 1u, flags, sleep_max,
				             C.sys && ax == 2 && (sgn < 0 ? peer_dn : peer_up));
		}
		if (!__syncthreads_and(ok)) break;
		LS_MARK(2);
		// ---- three upwind halo faces (sweep coordinate -1 on one axis) ---------------------------------------
		{
			// x face: (b, c) = 16 x 8 cells, threads 0..127
			const int b = tid % LS_TY, c = (tid / LS_TY) % LS_TZ;
			const int qx = dx > 0 ? x0 - 1 : x0 + wx, qy = y0 + (dy > 0 ? b : wy - 1 - b), qz = z0 + (dz > 0 ? c : wz - 1 - c);
			if (tid >= LS_LINES) {}
			else if (b < wy && c < wz && qx >= 0 && qx < d.nx) cp_async16(sm + SMI(-1, b, c), rec + ((size_t)qx + sY * (size_t)qy + sZ * (size_t)qz));
			else sm[SMI(-1, b, c)] = none;
		}
		{
			// y face: (a, c) = 16 x 8 cells, threads 128..255
			const int a = tid % LS_TX, c = (tid / LS_TX) % LS_TZ;
			const int qx = x0 + (dx > 0 ? a : wx - 1 - a), qy = dy > 0 ? y0 - 1 : y0 + wy, qz = z0 + (dz > 0 ? c : wz - 1 - c);
			if (tid < LS_LINES) {}
			else if (a < wx && c < wz && qy >= 0 && qy < d.ny) cp_async16(sm + SMI(a, -1, c), rec + ((size_t)qx + sY * (size_t)qy + sZ * (size_t)qz));
			else sm[SMI(a, -1, c)] = none;
		}
#pragma unroll
		for (int i = 0; i < LS_TX * LS_TY / LS_THREADS; i++) {
			// z face: (a, b) = 16 x 16 cells, one per thread
			const int e = tid + LS_THREADS * i, a = e % LS_TX, b = e / LS_TX;
			const int qx = x0 + (dx > 0 ? a : wx - 1 - a), qy = y0 + (dy > 0 ? b : wy - 1 - b), qz = dz > 0 ? z0 - 1 : z0 + wz;
			if (a < wx && b < wy && qz >= 0 && qz < d.nz) cp_async16(sm + SMI(a, b, -1), rec + ((size_t)qx + sY * (size_t)qy + sZ * (size_t)qz));
			else sm[SMI(a, b, -1)] = none;
		}
		cp_async_wait_all();
		__syncthreads();
		LS_MARK(3);
		// ---- squared distance of every own cell to its current record, and the quiet-tile test ---------------------------
		// Fully parallel, x fastest (conflict-free): if no cell would accept any of its three upwind candidates as
		// they stand BEFORE the sweep (q2 < m2 is necessary for acceptance: sqrt and "- radius" are monotonic),
		// then by induction along the sweep order the Gauss-Seidel pass changes nothing and the march is skipped.
		// The hyperplanes a+b+c that hold such a cell are collected in a bit mask: the march only runs a hyperplane that
		// holds one, or that follows a hyperplane in which a record changed.
		unsigned long long planes = 0;
		if (tid == 0) { s_planes[0] = 0u; s_planes[1] = 0u; }
		__syncthreads();
		{
			const int a = tid % LS_TX, b = tid / LS_TX;   // 16 x 16 threads, looping over c
			const float cxf = (float)(dx > 0 ? x0 + a : x0 + wx - 1 - a), cyf = (float)(dy > 0 ? y0 + b : y0 + wy - 1 - b);
#pragma unroll
			for (int c = 0; c < LS_TZ; c++) {
				const float czf = (float)(dz > 0 ? z0 + c : z0 + wz - 1 - c);
				const int i = SMI(a, b, c);
				const float4 cur = sm[i], X = sm[i - 1], Y = sm[i - LS_PX], Z = sm[i - LS_PX * LS_PY];
				// no record: infinitely far (any valid candidate wins); the same record in a neighbour has the same d^2 and never wins
				const float m2 = __float_as_int(cur.w) >= 0 ? ls_d2(cxf, cyf, czf, cur) : LS_INF;
				m2s[i] = m2;
				const bool wX = __float_as_int(X.w) > 0 && ls_d2(cxf, cyf, czf, X) < m2;
				const bool wY = __float_as_int(Y.w) > 0 && ls_d2(cxf, cyf, czf, Y) < m2;
				const bool wZ = __float_as_int(Z.w) > 0 && ls_d2(cxf, cyf, czf, Z) < m2;
				if (a < wx && b < wy && c < wz && (wX || wY || wZ)) planes |= 1ull << (a + b + c);
			}
			const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)planes), hi = __reduce_or_sync(0xffffffffu, (unsigned)(planes >> 32));
			if ((tid & 31) == 0) { if (lo) atomicOr(&s_planes[0], lo); if (hi) atomicOr(&s_planes[1], hi); }
		}
		__syncthreads();   // (also orders the m2s stores before the march)
		planes = (unsigned long long)s_planes[0] | ((unsigned long long)s_planes[1] << 32);
		const bool quiet = planes == 0;
#ifdef FS3_LS_TRACE
		if (tid == 0 && quiet) g_ls_trace[(size_t)t * 8 + 6] = 1;
#endif
		// ---- the cell hyperplanes; thread = line (b, c), marching along a ------------------------------------
		// clsInner on the upwind x-, y-, z-neighbour in that order (S3D.cpp:310-315 and siblings): index <= 0 is never
		// propagated (:242); a cell without a record takes any candidate and strict `dist < phi` keeps the incumbent
		// (:248) -- with "no record" and "no candidate" at distance +inf both are one comparison.
		// Fast path, branch-free, in squared distances: q2 >= m2 can never win (sqrt and "- radius" are monotonic)
		// and q2 < m2 by more than 2^-21 relative with sqrt(m2) >= 0.5 always wins (the sqrt values differ by >= 1 ulp
		// and sqrt(m2) - 1 is exact).  A comparison that falls between the two marks the step ambiguous and the
		// whole warp repeats it with the reference's full expression (warp-uniform branch, rare).
		// Idle threads compute garbage on a clamped cell of their own line and store nothing; only the 128 line
		// threads take part (named barrier 1).
		if (!quiet && tid < LS_LINES) {
			const int b = tid % LS_TY, c = tid / LS_TY;
			const bool line_in = b < wy && c < wz;
			const int gy = y0 + (dy > 0 ? b : wy - 1 - b), gz = z0 + (dz > 0 ? c : wz - 1 - c);
			const float cyf = (float)gy, czf = (float)gz;
			const int li = SMI(0, b, c);
			const int gx0 = dx > 0 ? x0 : x0 + wx - 1;
			float4* const grow = rec + (sY * (size_t)gy + sZ * (size_t)gz);
			uint32_t chg = 0;
			const int nsteps = wx + wy + wz - 2;
			for (int s = __ffsll((long long)planes) - 1; s < nsteps; s++) {
				const int a = s - b - c;
				const bool act = line_in && (unsigned)a < (unsigned)wx;
				const int ac = min(max(a, 0), LS_TX - 1);
				const float4* q = sm + li + ac;
				const float4 mine0 = q[0], X = q[-1], Y = q[-LS_PX], Z = q[-LS_PX * LS_PY];
				const float m20 = m2s[li + ac];
				const float cxf = (float)(gx0 + dx * ac);
				// a candidate with index <= 0 is never propagated (S3D.cpp:242): infinitely far
				const float x2 = __float_as_int(X.w) > 0 ? ls_d2(cxf, cyf, czf, X) : LS_INF;
				const float y2 = __float_as_int(Y.w) > 0 ? ls_d2(cxf, cyf, czf, Y) : LS_INF;
				const float z2 = __float_as_int(Z.w) > 0 ? ls_d2(cxf, cyf, czf, Z) : LS_INF;
				float4 mine = mine0;
				float m2 = m20;
				bool changed = false, amb = false;
#define LS_FAST(Q, q2)                                                                                                        \
	{                                                                                                                        \
		const bool take = m2 >= 0.25f && q2 < __fmul_rn(m2, 0.99999905f);                                                    \
		amb |= !take && q2 < m2;                                                                                             \
		mine.x = take ? Q.x : mine.x; mine.y = take ? Q.y : mine.y; mine.z = take ? Q.z : mine.z; mine.w = take ? Q.w : mine.w; \
		m2 = take ? q2 : m2; changed |= take;                                                                                \
	}
				LS_FAST(X, x2)
				LS_FAST(Y, y2)
				LS_FAST(Z, z2)
#undef LS_FAST
				if (__any_sync(0xffffffffu, amb && act)) {
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
				// global write-back is deferred past the loop: a store in flight would put an L2 round trip into every step
				if (act && changed) { sm[li + a] = mine; m2s[li + a] = m2; chg |= 1u << a; }
				// barrier among the 128 line threads + "did any record change in this hyperplane"
				int any;
				asm volatile("{\n .reg .pred p, q;\n setp.ne.u32 q, %1, 0;\n barrier.red.or.pred p, 1, %2, q;\n selp.u32 %0, 1, 0, p;\n}\n"
				             : "=r"(any) : "r"((unsigned)(act && changed)), "n"(LS_LINES) : "memory");
				if (!any) {
					// nothing changed: the next hyperplane that can is the next one holding a candidate cell
					const unsigned long long rest = s + 1 < 64 ? planes >> (s + 1) : 0ull;
					if (!rest) break;
					s += __ffsll((long long)rest) - 1;
				}
			}
#pragma unroll
			for (int a = 0; a < LS_TX; a++)
				if ((chg >> a) & 1u) __stcg(grow + (gx0 + dx * a), sm[li + a]);
		}
		if (m == 7) {
			// cells never reached keep their previous phi, as in the reference where m_Phi is only written
			// together with closestParticles (S3D.cpp:248-251,296-299).  x fastest: coalesced stores.
			float* const phi = reinterpret_cast<float*>(C.base[q] + C.off_phi);
			int* const closest = reinterpret_cast<int*>(C.base[q] + C.off_closest);
			// the projection of a neighbouring rank reads phi one plane into this slab: boundary tiles store that plane there too
			float* const phi_dn = peer_dn ? reinterpret_cast<float*>(C.base[q - 1] + C.off_phi) : nullptr;
			float* const phi_up = peer_up ? reinterpret_cast<float*>(C.base[q + 1] + C.off_phi) : nullptr;
			__syncthreads();
#pragma unroll
			for (int i = 0; i < LS_TX * LS_TY * LS_TZ / LS_THREADS; i++) {
				const int e = tid + LS_THREADS * i, px = e % LS_TX, py = (e / LS_TX) % LS_TY, pz = e / (LS_TX * LS_TY);
				if (px < wx && py < wy && pz < wz) {
					const int a = dx < 0 ? wx - 1 - px : px, b = dy < 0 ? wy - 1 - py : py, c = dz < 0 ? wz - 1 - pz : pz;
					const int id = __float_as_int(sm[SMI(a, b, c)].w);
					const size_t gi = (size_t)(x0 + px) + sY * (size_t)(y0 + py) + sZ * (size_t)(z0 + pz);
					closest[gi] = id;
					if (id >= 0) {
						const float ph = __fsub_rn(sqrtf(m2s[SMI(a, b, c)]), radius);
						phi[gi] = ph;
						if (phi_dn && pz == 0) phi_dn[gi] = ph;
						if (phi_up && pz == wz - 1) phi_up[gi] = ph;
					}
				}
			}
		}
		// ---- publish: the CTA's stores, then the epoch (release is cumulative over the barrier) ------------
		LS_MARK(4);
		__syncthreads();
		if (tid == 0) { __threadfence(); st_release_u32(mydone, E); }
		// ---- slab boundary: the plane the neighbouring rank reads as its upwind halo in this sweep direction goes into that
		// rank's copy of the records, then this tile's epoch into its copy of the epochs (every sweep: its tiles also wait for
		// "all six neighbours have finished the previous sweep") ---------------------------------------------------------------
		if (peer_dn || peer_up) {
			const int px = tid % LS_TX, py = tid / LS_TX;   // 16 x 16 cells of a plane, one per thread
			if (px < wx && py < wy) {
				const int a = dx < 0 ? wx - 1 - px : px, b = dy < 0 ? wy - 1 - py : py;
				// (the physical top plane is the last one of an upward sweep, the physical bottom plane the last one of a downward sweep)
				if (peer_up && dz > 0)     // read by the rank above in upward sweeps
					reinterpret_cast<float4*>(C.base[q + 1] + C.off_rec)[(size_t)(x0 + px) + sY * (size_t)(y0 + py) + sZ * (size_t)(z0 + wz - 1)] = sm[SMI(a, b, wz - 1)];
				if (peer_dn && dz < 0)     // read by the rank below in downward sweeps
					reinterpret_cast<float4*>(C.base[q - 1] + C.off_rec)[(size_t)(x0 + px) + sY * (size_t)(y0 + py) + sZ * (size_t)z0] = sm[SMI(a, b, wz - 1)];
			}
			__threadfence_system();
			__syncthreads();
			if (tid == 0) {
				if (peer_up) fs3_st_release_sys(reinterpret_cast<uint32_t*>(C.base[q + 1] + C.off_done) + tile_idx, E);
				if (peer_dn) fs3_st_release_sys(reinterpret_cast<uint32_t*>(C.base[q - 1] + C.off_done) + tile_idx, E);
			}
		}
		LS_MARK(5);
	}
#undef SMI
}

// seed of the cells of this rank's slab (world == 1: all cells)
int fs3i_levelset_seed(fs3_sim* s, cudaStream_t st) {
	GridDims d{s->nx, s->ny, s->nz};
	if (!st) st = s->stream;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SEED], st);
	const size_t plane = (size_t)s->nx * s->ny, c0 = plane * s->z0, c1 = plane * s->z1;
	k_ls_seed<<<cdiv((long long)(c1 - c0), 256), 256, 0, st>>>(s->pos[s->cur], s->cell_start, d, s->prm.particle_radius, s->ls_rec, c0, c1);
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

// Sharded: the boundary tiles of the neighbouring ranks have finished the last sweep, i.e. their phi planes have landed here.
__global__ void k_ls_wait_layers(const uint32_t* done, int ntiles_layer, size_t lo_first, size_t hi_first, int has_lo, int has_hi, uint32_t need, int* flags, int nospin) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= ntiles_layer) return;
	if (has_lo) fs3_flag_wait(done + lo_first + t, need, flags, nospin);
	if (has_hi) fs3_flag_wait(done + hi_first + t, need, flags, nospin);
}

// The 8 sweeps.  world == 1: the whole grid.  Multi-process sharded handle: the items of this rank's tiles; the ranks' kernels
// run at the same time on their GPUs and hand boundary planes and epochs to each other (see LsCtx).  Lockstep group: rank 0
// launches ONE kernel over the items of all ranks (kernels of different ranks on one device could not wait for each other),
// the other ranks only advance their epoch counters.
int fs3i_levelset_sweeps(fs3_sim* s, cudaStream_t st) {
	GridDims d{s->nx, s->ny, s->nz};
	float radius = s->prm.particle_radius;
	if (!st) st = s->stream;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SWEEP], st);
	const int slab = s->nz / s->world, lps = cdiv(slab, LS_TZ);
	int ntx = cdiv(d.nx, LS_TX), nty = cdiv(d.ny, LS_TY), ntz = s->world * lps;
	if (s->ls_nitems > 0) {
		LsCtx C;
		for (int r = 0; r < FS3_MAX_WORLD; r++) C.base[r] = s->world > 1 && r < s->world ? s->peer_base[r] : nullptr;
		const char* origin = s->world > 1 ? s->arena : nullptr;
		C.off_rec = (size_t)((char*)s->ls_rec - origin); C.off_done = (size_t)((char*)s->ls_done - origin);
		C.off_phi = (size_t)((char*)s->phi - origin); C.off_closest = (size_t)((char*)s->closest - origin);
		C.world = s->world; C.slab = slab; C.lps = lps; C.sys = (s->world > 1 && !s->lockstep) ? 1 : 0;
		FS3_CUDA(cudaMemsetAsync(s->ls_ticket, 0, sizeof(uint32_t), st));
#ifdef FS3_LS_TRACE
		static unsigned long long* trace_dev = nullptr;
		const size_t trace_n = (size_t)s->ls_nitems * 8;
		if (!trace_dev) { FS3_CUDA(cudaMalloc((void**)&trace_dev, trace_n * 8)); FS3_CUDA(cudaMemcpyToSymbol(g_ls_trace, &trace_dev, sizeof(trace_dev))); }
		FS3_CUDA(cudaMemsetAsync(trace_dev, 0, trace_n * 8, st));
#endif
		k_ls_sweep_dataflow<<<s->ls_blocks, LS_THREADS, LS_SMEM_BYTES, st>>>(C, d, radius, s->ls_order, s->ls_ticket, (uint32_t)s->ls_nitems, ntx, nty, ntz, s->ls_epoch,
		                                                                      s->dev_flags, (unsigned)s->ls_sleep_ns);
		FS3_LAUNCH_CHECK();
#ifdef FS3_LS_TRACE
		if (const char* path = getenv("FS3_LS_TRACE_FILE")) {
			std::vector<unsigned long long> h(trace_n);
			std::vector<uint32_t> ord((size_t)s->ls_nitems);
			FS3_CUDA(cudaStreamSynchronize(st));
			FS3_CUDA(cudaMemcpy(h.data(), trace_dev, trace_n * 8, cudaMemcpyDeviceToHost));
			FS3_CUDA(cudaMemcpy(ord.data(), s->ls_order, ord.size() * 4, cudaMemcpyDeviceToHost));
			if (FILE* fp = fopen(path, "wb")) { fwrite(h.data(), 8, trace_n, fp); fwrite(ord.data(), 4, ord.size(), fp); fclose(fp); }
		}
#endif
	}
	s->ls_epoch += 8;
	if (s->world > 1 && !s->lockstep) {
		// phi one plane beyond the slab (read by the projection) is stored by the neighbours' boundary tiles in their last sweep
		const int per_layer = ntx * nty;
		const bool lo = s->rank > 0, hi = s->rank < s->world - 1;
		k_ls_wait_layers<<<cdiv(per_layer, 128), 128, 0, st>>>(s->ls_done, per_layer, (size_t)per_layer * (size_t)(lo ? s->rank * lps - 1 : 0),
		                                                       (size_t)per_layer * (size_t)(hi ? (s->rank + 1) * lps : 0), lo ? 1 : 0, hi ? 1 : 0, s->ls_epoch, s->dev_flags, 0);
		FS3_LAUNCH_CHECK();
	}
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

// Work-item table, completion epochs and the persistent grid size (called once from fs3_create).
// Items (sweep m, tile) are sorted by their level in the dependency DAG -- the earliest "time" at which the
// item could run if every item took one unit -- which is a topological order (every dependency has a
// strictly smaller level), so ticket order is deadlock-free, and it lets sweep m+1 start behind sweep m as
// soon as the DAG allows instead of after it.
int fs3i_levelset_init(fs3_sim* s) {
	// tile layers are aligned to the z-slabs (world == 1: one slab), so that every tile belongs to exactly one rank
	const int slab = s->nz / s->world, lps = cdiv(slab, LS_TZ);
	const int ntx = cdiv(s->nx, LS_TX), nty = cdiv(s->ny, LS_TY), ntz = s->world * lps;
	if (ntx > 511 || nty > 511 || ntz > 511) { s->err = "grid too large for the level-set tile table"; return FS3_ERR_INVALID; }
	const int nt = ntx * nty * ntz;
	auto tid = [&](int x, int y, int z) { return x + ntx * (y + nty * z); };
	std::vector<int> prev(nt, -1), cur(nt, 0);
	std::vector<std::pair<int, uint32_t>> items;
	items.reserve((size_t)8 * nt);
	int maxlevel = 0;
	for (int m = 0; m < 8; m++) {
		const int dx = (m & 1) ? -1 : 1, dy = (m & 2) ? -1 : 1, dz = (m & 4) ? -1 : 1;
		for (int kz = 0; kz < ntz; kz++)
			for (int ky = 0; ky < nty; ky++)
				for (int kx = 0; kx < ntx; kx++) {   // sweep-space order: upwind neighbours are visited first
					const int x = dx > 0 ? kx : ntx - 1 - kx, y = dy > 0 ? ky : nty - 1 - ky, z = dz > 0 ? kz : ntz - 1 - kz;
					int L = 0;
					if (kx > 0) L = std::max(L, cur[tid(x - dx, y, z)] + 1);
					if (ky > 0) L = std::max(L, cur[tid(x, y - dy, z)] + 1);
					if (kz > 0) L = std::max(L, cur[tid(x, y, z - dz)] + 1);
					if (m > 0) {
						L = std::max(L, prev[tid(x, y, z)] + 1);
						if (x > 0) L = std::max(L, prev[tid(x - 1, y, z)] + 1);
						if (x < ntx - 1) L = std::max(L, prev[tid(x + 1, y, z)] + 1);
						if (y > 0) L = std::max(L, prev[tid(x, y - 1, z)] + 1);
						if (y < nty - 1) L = std::max(L, prev[tid(x, y + 1, z)] + 1);
						if (z > 0) L = std::max(L, prev[tid(x, y, z - 1)] + 1);
						if (z < ntz - 1) L = std::max(L, prev[tid(x, y, z + 1)] + 1);
					}
					cur[tid(x, y, z)] = L;
					maxlevel = std::max(maxlevel, L);
					items.emplace_back(L, (uint32_t)x | ((uint32_t)y << 9) | ((uint32_t)z << 18) | ((uint32_t)m << 27));
				}
		prev = cur;
	}
	std::stable_sort(items.begin(), items.end(), [](const std::pair<int, uint32_t>& a, const std::pair<int, uint32_t>& b) { return a.first < b.first; });
	// The items this handle launches, in the order of the GLOBAL dependency DAG: all of them on one GPU and on rank 0 of a
	// lockstep group, the tiles of the own slab on a multi-process rank (every rank then draws its items in an order in which
	// each dependency, local or on another rank, precedes its dependants: no deadlock), none on the other lockstep ranks.
	std::vector<uint32_t> ord;
	ord.reserve(items.size());
	for (size_t i = 0; i < items.size(); i++) {
		const int owner = (int)((items[i].second >> 18) & 511u) / lps;
		if (s->world == 1 || (s->lockstep ? s->rank == 0 : owner == s->rank)) ord.push_back(items[i].second);
	}
	s->ls_nitems = (long long)ord.size();
	if (ord.empty()) ord.push_back(0u);
	int rc;
	if ((rc = fs3i_alloc(s, (void**)&s->ls_order, sizeof(uint32_t) * 8 * (size_t)nt))) return rc;   // (same size on every rank: one arena layout)
	if ((rc = fs3i_alloc(s, (void**)&s->ls_done, sizeof(uint32_t) * nt))) return rc;
	if ((rc = fs3i_alloc(s, (void**)&s->ls_ticket, sizeof(uint32_t)))) return rc;
	FS3_CUDA(cudaMemcpyAsync(s->ls_order, ord.data(), sizeof(uint32_t) * ord.size(), cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaMemsetAsync(s->ls_done, 0, sizeof(uint32_t) * nt, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	s->ls_ntiles = nt;
	s->ls_levels = maxlevel + 1;
	s->ls_epoch = 0;
	int per_sm = 0, sms = 0;
	const size_t smem = LS_SMEM_BYTES;
	FS3_CUDA(cudaFuncSetAttribute(k_ls_sweep_dataflow, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
	FS3_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ls_sweep_dataflow, LS_THREADS, smem));
	FS3_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
	// Resident CTAs beyond what the DAG can keep busy only add pollers.
	const int occ = std::max(1, per_sm);
	// The sweep is bound by its critical path, not by throughput, and it runs beside the P2G gather and the extrapolation
	// (fs3_step), which are issue-bound: resident sweep CTAs beyond what the DAG keeps busy only add pollers and take issue slots
	// from that chain (measured round 1, one GPU: 19.55 -> 18.9 ms per step with one CTA per SM instead of two).  There is also
	// a hard reason for ONE CTA per SM: a sweep CTA holds 256 x 80 registers until the last item is done, and a P2G CTA needs
	// 512 x 64 -- beside two resident sweep CTAs no P2G CTA fits on the SM at all, and the gather would run after the sweeps
	// instead of beside them.  A sharded rank runs 1 / world of the tiles per dependency level, so one CTA per SM is plenty.
	per_sm = std::min(occ, 1);
	s->ls_sleep_ns = 256;
	long long blocks = (long long)std::max(1, per_sm) * sms;
	s->ls_blocks = (int)std::max<long long>(1, std::min<long long>(blocks, s->ls_nitems));
	return FS3_OK;
}
