// Particle-side kernels: reset (closed-form LCG skip-ahead), RK3 advection, counting sort by cell,
// PIC/FLIP grid-to-particle blend.  sm_100a; compiled with -fmad=false (the reference build never
// contracts a*b+c).  All kernels are particle-parallel, 128-bit loads/stores on the
// float4 SoA arrays; particles are kept physically sorted by cell so that the grid gathers of one
// warp hit a handful of neighbouring cache lines.
#include "fs3_internal.cuh"
#include <algorithm>

// ------------------------------------------------------------------------------------------------
// ResetSimulation (S3D.cpp:41-99).  std::default_random_engine(0) is minstd_rand0 in libstdc++:
// x_k = 16807^k mod (2^31-1) with x_0 = 1, one draw per value, so draw k is computable independently.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t lcg_state(unsigned long long k) {
	const unsigned long long M = 2147483647ull;
	unsigned long long r = 1, a = 16807ull;
	while (k) {
		if (k & 1) r = (r * a) % M;
		a = (a * a) % M;
		k >>= 1;
	}
	return (uint32_t)r;
}
// uniform_real_distribution<float>(-0.25f, 0.25f) of draw k >= 1 (libstdc++ generate_canonical<float,24>)
__device__ __forceinline__ float lcg_uniform(unsigned long long k) {
	uint32_t x = lcg_state(k);
	float r = (float)(x - 1u) / 2147483648.0f;
	if (r >= 1.0f) r = 0.99999994f;  // nextafterf(1,0)
	return __fadd_rn(__fmul_rn(r, 0.5f), -0.25f);
}

__global__ void k_reset_grids(float* __restrict__ U, float* __restrict__ V, float* __restrict__ W, size_t nU, size_t nV, size_t nW) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	size_t tot = nU + nV + nW;
	for (; q < tot; q += (size_t)gridDim.x * blockDim.x) {
		float r = lcg_uniform(q + 1);
		if (q < nU) U[q] = r;
		else if (q < nU + nV) V[q - nU] = r;
		else W[q - nU - nV] = r;
	}
}

// creation ids [j0, j0+np) are written to pos[0..np) (a z-slab of the block generates a contiguous id range)
__global__ void k_reset_particles(float4* __restrict__ pos, float4* __restrict__ vel, long long np, long long j0, GridDims d, float cpm,
                                  const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W,
                                  unsigned long long draw0) {
	long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (j >= np) return;
	pos -= j0; vel -= j0;   // slot of creation id j is j - j0
	j += j0;
	int c = (int)(j & 1), b = (int)((j >> 1) & 1), a = (int)((j >> 2) & 1);
	long long cellj = j >> 3;
	int xc = d.nx - 1 - d.nx / 2, yc = d.ny - 2;
	int x = d.nx / 2 + (int)(cellj % xc);
	int y = 1 + (int)((cellj / xc) % yc);
	int z = 1 + (int)(cellj / ((long long)xc * yc));
	float dd = 0.5f / cpm;
	unsigned long long k = draw0 + 3ull * (unsigned long long)j;
	float m1 = __fadd_rn(__fadd_rn(((float)x - 0.25f) / cpm, __fmul_rn((float)a, dd)), lcg_uniform(k) / cpm);
	float m2 = __fadd_rn(__fadd_rn(((float)y - 0.25f) / cpm, __fmul_rn((float)b, dd)), lcg_uniform(k + 1) / cpm);
	float m3 = __fadd_rn(__fadd_rn(((float)z - 0.25f) / cpm, __fmul_rn((float)c, dd)), lcg_uniform(k + 2) / cpm);
	float3 u = fs3_interp(U, V, W, d, __fmul_rn((float)d.nx, m1), __fmul_rn((float)d.ny, m2), __fmul_rn((float)d.nz, m3));
	pos[j] = make_float4(m1, m2, m3, __int_as_float((int)j));
	vel[j] = make_float4(u.x, u.y, u.z, 0.0f);
}

int fs3i_reset_grids(fs3_sim* s) {
	size_t tot = s->nU + s->nV + s->nW;
	k_reset_grids<<<cdiv((long long)tot, 256 * 4), 256, 0, s->stream>>>(s->U, s->V, s->W, s->nU, s->nV, s->nW);
	FS3_LAUNCH_CHECK();
	s->cur = 0;
	s->sorted = false; s->counted = false;
	s->closest_valid = false;
	s->system_valid = false;
	return FS3_OK;
}

// Particles whose GENERATING cell lies in this rank's slab (world == 1: all of them); the generating cells of the dam-break
// block are z in [1, nz-1), z slowest, so a slab is a contiguous range of creation ids.  A jittered particle can round into
// the neighbouring slab: the sharded reset runs one migration afterwards.
int fs3i_reset_particles_slab(fs3_sim* s) {
	size_t tot = s->nU + s->nV + s->nW;
	const long long per_layer = 8LL * (s->ny - 2) * (s->nx - 1 - s->nx / 2);
	for (int r = 0; r < s->world; r++) {
		int zs = s->nz / s->world, a = std::max(r * zs, 1), b = std::min((r + 1) * zs, s->nz - 1);
		s->np_all[r] = per_layer * std::max(0, b - a);
	}
	const int zlo = std::max(s->z0, 1), zhi = std::min(s->z1, s->nz - 1);
	const long long np = per_layer * std::max(0, zhi - zlo), j0 = per_layer * (zlo - 1);
	if (np > s->cap) { s->err = "particle capacity of the slab exceeded at reset"; return FS3_ERR_STATE; }
	s->np = s->np_in = np;
	GridDims d{s->nx, s->ny, s->nz};
	if (np > 0) {
		k_reset_particles<<<cdiv(np, 256), 256, 0, s->stream>>>(s->pos[0] + s->poff, s->vel[0] + s->poff, np, j0, d, s->cpm, s->U, s->V, s->W,
		                                                         (unsigned long long)tot + 1ull);
		FS3_LAUNCH_CHECK();
	}
	return FS3_OK;
}

int fs3i_reset(fs3_sim* s) {
	int rc = fs3i_reset_grids(s);
	if (rc) return rc;
	return fs3i_reset_particles_slab(s);
}

// ------------------------------------------------------------------------------------------------
// Advect (S3D.cpp:190-232): Ralston RK3 through the grid, clamp into the box.
// ------------------------------------------------------------------------------------------------
// Emigrant bookkeeping of a sharded handle: destination rank = slab of the particle's cell, (int)roundf(n*Z) (S3D.cpp:288).
struct MigArgs {
	uint32_t* ctl;       // [0,world) per-destination counts, [world,2world) cursors, [2world] number of emigrants
	uint32_t* leaver;    // their indices in the live range
	uint32_t cap;
	int rank, world, slab, nz;
	// index-space z range inside which every tap of InterpolateMACCell falls into planes of the projected field this rank holds
	// (its slab + the halo its neighbours push, fs3i_velocity_halo); flags[4] reports a Runge-Kutta stage outside it
	float zmin_ok, zmax_ok;
	int* flags;
};
__device__ __forceinline__ int mig_owner(const MigArgs& m, float z) {
	int cz = min(max(fs3_cell_axis(z, (float)m.nz), 0), m.nz - 1);
	return cz / m.slab;
}
__device__ __forceinline__ void mig_note(const MigArgs& m, long long i, float z) {
	int dest = mig_owner(m, z);
	if (dest != m.rank) {
		atomicAdd(m.ctl + dest, 1u);
		uint32_t k = atomicAdd(m.ctl + 2 * m.world, 1u);
		if (k < m.cap) m.leaver[k] = (uint32_t)i;
	}
}

// Half-cell key of a particle for the counting sort (see "Binning" below) and its count; shared by Advect (single GPU: the
// sort's counting pass is fused into it) and k_bin_count.
struct BinArgs {
	uint32_t* cellid;    // per particle: half-cell key
	uint32_t* arrival;   // per particle: arrival order inside the half-cell
	uint32_t* count;     // per half-cell
	int* flags;
	int zlo, zhi;
};
__device__ __forceinline__ void bin_note(const BinArgs& B, long long i, float4 p, GridDims d) {
	int cx, cy, cz;
	int hx = fs3_halfcell_axis(p.x, (float)d.nx, cx), hy = fs3_halfcell_axis(p.y, (float)d.ny, cy), hz = fs3_halfcell_axis(p.z, (float)d.nz, cz);
	if (cx < 0 || cx >= d.nx || cy < 0 || cy >= d.ny || cz < B.zlo || cz >= B.zhi) {
		B.flags[0] = 1;  // outside the box (or the slab): the reference skips such particles; Advect's clamp makes this unreachable
		hx = min(max(hx, 0), 2 * d.nx - 1); hy = min(max(hy, 0), 2 * d.ny - 1); hz = min(max(hz, 2 * B.zlo), 2 * B.zhi - 1);
	}
	uint32_t c = (uint32_t)hx + (uint32_t)(2 * d.nx) * ((uint32_t)hy + (uint32_t)(2 * d.ny) * (uint32_t)hz);
	B.cellid[i] = c;
	B.arrival[i] = atomicAdd(B.count + c, 1u);   // arrival order inside the half-cell: a unique provisional slot (the rank by creation id follows)
}

// Clamp box of Advect, [(-0.5+0.1)/n, 1+(-0.5-0.1)/n] per axis (S3D.cpp:226-230): six IEEE float divisions that depend on the
// grid only, so the host does them once per launch instead of every thread (same single-precision operations, same results).
struct AdvectBox {
	float lo[3], hi[3];
};
static AdvectBox advect_box(const fs3_sim* s) {
	AdvectBox b;
	const float eps = 0.1f, n[3] = {(float)s->nx, (float)s->ny, (float)s->nz};
	for (int a = 0; a < 3; a++) {
		volatile float lo = (-0.5f + eps) / n[a], q = (-0.5f - eps) / n[a];   // volatile: keep every intermediate a rounded float
		volatile float hi = 1.0f + q;
		b.lo[a] = lo; b.hi[a] = hi;
	}
	return b;
}

template <bool SHARD>
__global__ void __launch_bounds__(256) k_advect(float4* __restrict__ pos, long long np, GridDims d, float dt,
                                                const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W, MigArgs mig,
                                                BinArgs bin, AdvectBox box) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i];
	if (__float_as_int(p.w) < 0) { if (!SHARD) bin.cellid[i] = 0xffffffffu; return; }   // dead slot (emigrated)
	float fx = (float)d.nx, fy = (float)d.ny, fz = (float)d.nz;
	const float z1c = __fmul_rn(fz, p.z);
	float3 k1 = fs3_interp(U, V, W, d, __fmul_rn(fx, p.x), __fmul_rn(fy, p.y), z1c);
	float h = __fmul_rn(0.5f, dt);
	const float z2c = __fmul_rn(fz, __fadd_rn(p.z, __fmul_rn(h, k1.z)));
	float3 k2 = fs3_interp(U, V, W, d, __fmul_rn(fx, __fadd_rn(p.x, __fmul_rn(h, k1.x))), __fmul_rn(fy, __fadd_rn(p.y, __fmul_rn(h, k1.y))), z2c);
	float t = __fmul_rn(0.75f, dt);
	const float z3c = __fmul_rn(fz, __fadd_rn(p.z, __fmul_rn(t, k2.z)));
	float3 k3 = fs3_interp(U, V, W, d, __fmul_rn(fx, __fadd_rn(p.x, __fmul_rn(t, k2.x))), __fmul_rn(fy, __fadd_rn(p.y, __fmul_rn(t, k2.y))), z3c);
	if (SHARD && (fminf(z1c, fminf(z2c, z3c)) < mig.zmin_ok || fmaxf(z1c, fmaxf(z2c, z3c)) > mig.zmax_ok)) mig.flags[4] = 1;
	const float c1 = 2.0f / 9.0f, c2 = 3.0f / 9.0f, c3 = 4.0f / 9.0f;
	float vx = __fadd_rn(__fadd_rn(__fmul_rn(c1, k1.x), __fmul_rn(c2, k2.x)), __fmul_rn(c3, k3.x));
	float vy = __fadd_rn(__fadd_rn(__fmul_rn(c1, k1.y), __fmul_rn(c2, k2.y)), __fmul_rn(c3, k3.y));
	float vz = __fadd_rn(__fadd_rn(__fmul_rn(c1, k1.z), __fmul_rn(c2, k2.z)), __fmul_rn(c3, k3.z));
	p.x = __fadd_rn(p.x, __fmul_rn(dt, vx));
	p.y = __fadd_rn(p.y, __fmul_rn(dt, vy));
	p.z = __fadd_rn(p.z, __fmul_rn(dt, vz));
	p.x = fs3_clamp(p.x, box.lo[0], box.hi[0]);
	p.y = fs3_clamp(p.y, box.lo[1], box.hi[1]);
	p.z = fs3_clamp(p.z, box.lo[2], box.hi[2]);
	pos[i] = p;
	if (SHARD) mig_note(mig, i, p.z);
	else bin_note(bin, i, p, d);
}

static MigArgs mig_args(fs3_sim* s) {
	MigArgs m;
	m.ctl = s->mig_ctl; m.leaver = s->leaver_idx; m.cap = (uint32_t)s->mig_cap;
	m.rank = s->rank; m.world = s->world; m.slab = s->nz / s->world; m.nz = s->nz;
	const int hv = fs3i_velocity_halo(s), wlo = std::max(0, s->z0 - hv), whi = std::min(s->nz, s->z1 + hv);
	m.zmin_ok = wlo == 0 ? -1e30f : (float)wlo + 1.0f;
	m.zmax_ok = whi == s->nz ? 1e30f : (float)whi - 1.5f;
	m.flags = s->dev_flags;
	return m;
}

int fs3i_advect(fs3_sim* s, float dt) {
	GridDims d{s->nx, s->ny, s->nz};
	s->sorted = false; s->counted = false;
	s->closest_valid = false;
	if (s->world > 1) {
		FS3_CUDA(cudaMemsetAsync(s->mig_ctl, 0, (2 * s->world + 1) * sizeof(uint32_t), s->stream));
		if (s->np_in > 0) {
			k_advect<true><<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->np_in, d, dt, s->U, s->V, s->W, mig_args(s), BinArgs{}, advect_box(s));
			FS3_LAUNCH_CHECK();
		}
		return FS3_OK;
	}
	if (s->np <= 0) return FS3_OK;
	// the counting pass of the sort that always follows is fused in (the new position is in registers)
	FS3_CUDA(cudaMemsetAsync(s->cell_count, 0, (s->nH + 1) * sizeof(uint32_t), s->stream));
	k_advect<false><<<cdiv(s->np, 256), 256, 0, s->stream>>>(s->pos[s->cur], s->np, d, dt, s->U, s->V, s->W, MigArgs{},
	                                                          BinArgs{s->cellid, s->cell_fill, s->cell_count, s->dev_flags, 0, s->nz}, advect_box(s));
	FS3_LAUNCH_CHECK();
	s->counted = true;
	return FS3_OK;
}

// Emigrants of the live range without moving anything (after reset / a particle upload).
__global__ void __launch_bounds__(256) k_mig_mark(const float4* __restrict__ pos, long long np, MigArgs mig) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i];
	if (__float_as_int(p.w) < 0) return;
	mig_note(mig, i, p.z);
}
int fs3i_mig_mark(fs3_sim* s) {
	FS3_CUDA(cudaMemsetAsync(s->mig_ctl, 0, (2 * s->world + 1) * sizeof(uint32_t), s->stream));
	if (s->np_in > 0) { k_mig_mark<<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->np_in, mig_args(s)); FS3_LAUNCH_CHECK(); }
	s->sorted = false; s->counted = false;
	return FS3_OK;
}

// Emigrants -> outgoing buffers grouped by destination rank; their slots become dead (id = -1).
__global__ void __launch_bounds__(256) k_mig_pack(float4* __restrict__ pos, const float4* __restrict__ vel, MigArgs mig, float4* __restrict__ opos,
                                                  float4* __restrict__ ovel, int* __restrict__ flags) {
	const uint32_t n = mig.ctl[2 * mig.world];
	if (n > mig.cap) { if (blockIdx.x == 0 && threadIdx.x == 0) flags[3] = 1; return; }
	for (uint32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
		const uint32_t i = mig.leaver[t];
		float4 p = pos[i];
		const int dest = mig_owner(mig, p.z);
		uint32_t base = 0;
		for (int r = 0; r < dest; r++) base += mig.ctl[r];
		const uint32_t slot = base + atomicAdd(mig.ctl + mig.world + dest, 1u);
		opos[slot] = p;
		ovel[slot] = vel[i];
		pos[i].w = __int_as_float(-1);
	}
}
int fs3i_mig_pack(fs3_sim* s) {
	k_mig_pack<<<std::max(1, std::min(cdiv(s->mig_cap, 256), s->sm_count * 4)), 256, 0, s->stream>>>(
	    s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff, mig_args(s), s->mig_pos, s->mig_vel, s->dev_flags);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// ------------------------------------------------------------------------------------------------
// Binning: counting sort of the particles by HALF-cell (cell = (int)roundf(n*X), S3D.cpp:286-288, refined by the
// particle's octant inside the cell; key on the (2n)^3 half-cell grid, x fastest -- see fs3_halfcell_axis).
// count -> exclusive scan -> scatter creation ids -> rank inside the cell by creation id -> gather.
// Ranking by creation id makes the order inside a cell the reference's particle order, so the result
// is deterministic (no dependence on atomic arrival order) and "lowest index wins" tie-breaks
// (S3D.cpp:296) fall out of a forward scan of the cell's segment.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_count(const float4* __restrict__ pos, long long np, GridDims d, BinArgs bin) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i];
	if (__float_as_int(p.w) < 0) { bin.cellid[i] = 0xffffffffu; return; }   // dead slot of an emigrant
	bin_note(bin, i, p, d);
}

// --- exclusive scan of uint32, tile = 512 threads x SCAN_I items (template parameter of the kernel) ----
#define SCAN_T 512

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* total) {
	__shared__ uint32_t wsum[SCAN_T / 32];
	int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
	uint32_t inc = v;
#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
		if (lane >= o) inc += t;
	}
	if (lane == 31) wsum[wid] = inc;
	__syncthreads();
	if (wid == 0) {
		uint32_t w = lane < SCAN_T / 32 ? wsum[lane] : 0;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
			if (lane >= o) w += t;
		}
		if (lane < SCAN_T / 32) wsum[lane] = w;
	}
	__syncthreads();
	uint32_t base = wid ? wsum[wid - 1] : 0;
	if (total) *total = wsum[SCAN_T / 32 - 1];
	__syncthreads();
	return base + inc - v;
}

// Single pass with decoupled look-back: tiles are handed out by a ticket (so a tile's predecessors are always running or
// done), every tile publishes its aggregate, then its inclusive prefix, as (flag << 32 | value) words, and resolves its
// own prefix by walking back over the predecessors' words.  The input is read once and the output written once.
#define SCAN_AGG 1ull
#define SCAN_INC 2ull
__device__ __forceinline__ unsigned long long scan_ld(const unsigned long long* p) {
	unsigned long long v;
	asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];\n" : "=l"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void scan_st(unsigned long long* p, unsigned long long v) { asm volatile("st.relaxed.gpu.global.u64 [%0], %1;\n" ::"l"(p), "l"(v) : "memory"); }

template <int SCAN_I>
__global__ void __launch_bounds__(SCAN_T) k_scan_lookback(const uint32_t* __restrict__ in, uint32_t* __restrict__ out, size_t n,
                                                          unsigned long long* __restrict__ state, uint32_t* __restrict__ ticket, uint32_t base0) {
	constexpr int SCAN_TILE = SCAN_T * SCAN_I;
	__shared__ uint32_t s_tile, s_prefix;
	if (threadIdx.x == 0) s_tile = atomicAdd(ticket, 1u);
	__syncthreads();
	const uint32_t tile = s_tile;
	const size_t base = (size_t)tile * SCAN_TILE + (size_t)threadIdx.x * SCAN_I;
	uint32_t loc[SCAN_I], v = 0;
	if (base + SCAN_I <= n && (reinterpret_cast<uintptr_t>(in + base) & 15) == 0) {
#pragma unroll
		for (int k = 0; k < SCAN_I; k += 4) {
			const uint4 a = *reinterpret_cast<const uint4*>(in + base + k);
			loc[k] = a.x; loc[k + 1] = a.y; loc[k + 2] = a.z; loc[k + 3] = a.w;
		}
	} else {
#pragma unroll
		for (int k = 0; k < SCAN_I; k++) loc[k] = base + k < n ? in[base + k] : 0;
	}
#pragma unroll
	for (int k = 0; k < SCAN_I; k++) v += loc[k];
	uint32_t tot;
	const uint32_t ex_in_tile = block_excl_scan(v, &tot);
	// look-back by the first warp, 32 predecessors per step: every lane waits for its predecessor's word, the nearest one
	// that already holds an inclusive prefix ends the walk
	if (threadIdx.x < 32) {
		const int lane = threadIdx.x;
		uint32_t prefix = 0;
		if (tile == 0) {
			if (lane == 0) scan_st(state, (SCAN_INC << 32) | tot);
		} else {
			if (lane == 0) scan_st(state + tile, (SCAN_AGG << 32) | tot);
			for (long long j = (long long)tile - 1;; j -= 32) {
				const long long idx = j - lane;
				unsigned long long w = SCAN_INC << 32;   // before the first tile: an inclusive prefix of 0
				if (idx >= 0) do { w = scan_ld(state + idx); } while ((w >> 32) == 0);
				const unsigned inc = __ballot_sync(0xffffffffu, (w >> 32) == SCAN_INC);
				const int first = inc ? __ffs(inc) - 1 : 31;
				uint32_t val = lane <= first ? (uint32_t)w : 0u;
#pragma unroll
				for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
				prefix += val;
				if (inc) break;
			}
			if (lane == 0) scan_st(state + tile, (SCAN_INC << 32) | (uint32_t)(prefix + tot));
		}
		if (lane == 0) s_prefix = prefix;
	}
	__syncthreads();
	uint32_t ex = ex_in_tile + s_prefix + base0;
	if (base + SCAN_I <= n && (reinterpret_cast<uintptr_t>(out + base) & 15) == 0) {
#pragma unroll
		for (int k = 0; k < SCAN_I; k += 4) {
			uint4 a;
			a.x = ex; ex += loc[k]; a.y = ex; ex += loc[k + 1]; a.z = ex; ex += loc[k + 2]; a.w = ex; ex += loc[k + 3];
			*reinterpret_cast<uint4*>(out + base + k) = a;
		}
	} else {
#pragma unroll
		for (int k = 0; k < SCAN_I; k++) { if (base + k < n) out[base + k] = ex; ex += loc[k]; }
	}
}

int fs3i_scan_u32(fs3_sim* s, const uint32_t* in, uint32_t* out, size_t n, uint32_t base0) {
	// items per thread: a tile's cost is a chain of latencies (ticket, load, block scan, look-back, store), so larger tiles
	// amortise it better -- BIN at 256^3 with 8 / 16 / 32 items: 1.80 / 1.67 / 1.65 ms.
	const int items = 32;
	int nb = cdiv((long long)n, SCAN_T * items);
	if ((size_t)nb > s->scan_block_cap) { s->err = "scan scratch too small"; return FS3_ERR_STATE; }
	// scratch: one 64-bit word per tile, then the ticket
	unsigned long long* state = reinterpret_cast<unsigned long long*>(s->scan_block);
	uint32_t* ticket = reinterpret_cast<uint32_t*>(state + s->scan_block_cap);
	FS3_CUDA(cudaMemsetAsync(s->scan_block, 0, ((size_t)nb + 0) * 8, s->stream));
	FS3_CUDA(cudaMemsetAsync(ticket, 0, 4, s->stream));
	k_scan_lookback<32><<<nb, SCAN_T, 0, s->stream>>>(in, out, n, state, ticket, base0);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

__global__ void __launch_bounds__(256) k_bin_scatter_ids(const float4* __restrict__ pos, const uint32_t* __restrict__ cellid, long long np,
                                                         const uint32_t* __restrict__ start, const uint32_t* __restrict__ arrival, int* __restrict__ tmp_id) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	uint32_t c = cellid[i];
	if (c == 0xffffffffu) return;
	uint32_t slot = start[c] + arrival[i];
	tmp_id[slot] = __float_as_int(pos[i].w);
}

__global__ void __launch_bounds__(256) k_bin_rank_gather(const float4* __restrict__ pin, const float4* __restrict__ vin,
                                                         const uint32_t* __restrict__ cellid, long long np, const uint32_t* __restrict__ start,
                                                         const int* __restrict__ tmp_id, float4* __restrict__ pout, float4* __restrict__ vout) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	uint32_t c = cellid[i];
	if (c == 0xffffffffu) return;
	float4 p = pin[i], v = vin[i];
	uint32_t b = start[c], e = start[c + 1];
	int id = __float_as_int(p.w);
	uint32_t rank = 0;
	for (uint32_t t = b; t < e; t++) rank += (__ldg(tmp_id + t) < id) ? 1u : 0u;
	pout[b + rank] = p;
	vout[b + rank] = v;
}

// Sorts the live range [poff, poff+np_in) of buffer `cur` (dead slots dropped) into [poff, poff+np) of the other buffer.
// A sharded handle only owns the half-cell keys of its slab, [K0, K1): counts, scan and cursors are restricted to them and
// cell_start holds ABSOLUTE slots (poff included), so that the neighbours' boundary layers can be spliced in below / above
// (fs3i_ghost_pull) and the gather kernels index pos/vel from the buffer base.
int fs3i_bin(fs3_sim* s) {
	if (s->sorted) return FS3_OK;
	const size_t hplane = (size_t)(2 * s->nx) * (2 * s->ny);
	const size_t K0 = hplane * (size_t)(2 * s->z0), K1 = hplane * (size_t)(2 * s->z1);
	if (!s->counted) FS3_CUDA(cudaMemsetAsync(s->cell_count + K0, 0, (K1 - K0 + 1) * sizeof(uint32_t), s->stream));
	if (s->np_in > 0) {
		GridDims d{s->nx, s->ny, s->nz};
		int nb = cdiv(s->np_in, 256);
		const float4 *pin = s->pos[s->cur] + s->poff, *vin = s->vel[s->cur] + s->poff;
		if (!s->counted) {
			k_bin_count<<<nb, 256, 0, s->stream>>>(pin, s->np_in, d, BinArgs{s->cellid, s->cell_fill, s->cell_count, s->dev_flags, s->z0, s->z1});
			FS3_LAUNCH_CHECK();
		}
		int rc = fs3i_scan_u32(s, s->cell_count + K0, s->cell_start + K0, K1 - K0 + 1, (uint32_t)s->poff);
		if (rc) return rc;
		k_bin_scatter_ids<<<nb, 256, 0, s->stream>>>(pin, s->cellid, s->np_in, s->cell_start, s->cell_fill, s->tmp_id); FS3_LAUNCH_CHECK();
		k_bin_rank_gather<<<nb, 256, 0, s->stream>>>(pin, vin, s->cellid, s->np_in, s->cell_start, s->tmp_id,
		                                              s->pos[s->cur ^ 1], s->vel[s->cur ^ 1]); FS3_LAUNCH_CHECK();
		s->cur ^= 1;
	} else {
		int rc = fs3i_scan_u32(s, s->cell_count + K0, s->cell_start + K0, K1 - K0 + 1, (uint32_t)s->poff);
		if (rc) return rc;
		// A sharded rank whose slab is empty flips its buffer like every other rank: the neighbours address this rank's live
		// buffer with THEIR OWN `cur` (immigrants in fs3i_mig_exchange, ghost layers in fs3i_ghost_pull), so `cur` must stay
		// identical on all ranks whatever each slab holds.
		if (s->world > 1) s->cur ^= 1;
	}
	s->np_in = s->np;
	s->sorted = true;
	s->counted = false;
	return FS3_OK;
}

// ------------------------------------------------------------------------------------------------
// Boundary conversions: AoS Particle3 in creation order <-> internal sorted SoA
// ------------------------------------------------------------------------------------------------
__global__ void k_unsort(const float4* __restrict__ pos, const float4* __restrict__ vel, long long np, fs3_particle* __restrict__ out) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i], v = vel[i];
	if (__float_as_int(p.w) < 0) return;
	fs3_particle q{p.x, p.y, p.z, v.x, v.y, v.z};
	out[__float_as_int(p.w)] = q;
}
__global__ void k_load(const fs3_particle* __restrict__ in, long long np, float4* __restrict__ pos, float4* __restrict__ vel) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	fs3_particle q = in[i];
	pos[i] = make_float4(q.X, q.Y, q.Z, __int_as_float((int)i));
	vel[i] = make_float4(q.uX, q.uY, q.uZ, 0.0f);
}
__global__ void k_cell_indices(const float4* __restrict__ pos, long long np, GridDims d, int* __restrict__ out) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i];
	if (__float_as_int(p.w) < 0) return;
	int cx = fs3_cell_axis(p.x, (float)d.nx), cy = fs3_cell_axis(p.y, (float)d.ny), cz = fs3_cell_axis(p.z, (float)d.nz);
	bool in = cx >= 0 && cx < d.nx && cy >= 0 && cy < d.ny && cz >= 0 && cz < d.nz;
	out[__float_as_int(p.w)] = in ? cx + d.nx * (cy + d.ny * cz) : -1;
}

int fs3i_unsort_particles(fs3_sim* s, fs3_particle* dev_out) {
	if (s->np_in <= 0) return FS3_OK;
	k_unsort<<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff, s->np_in, dev_out); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_load_particles(fs3_sim* s, const fs3_particle* dev_in, long long count) {
	s->np = s->np_in = count; s->cur = 0; s->sorted = false; s->counted = false; s->closest_valid = false;
	if (count <= 0) return FS3_OK;
	k_load<<<cdiv(count, 256), 256, 0, s->stream>>>(dev_in, count, s->pos[0], s->vel[0]); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
// Sharded upload: every rank is handed the SAME full array and keeps the particles whose cell lies in its slab
// (creation id = array index).  The per-rank counts come out of the same pass, so no communication is needed.
__global__ void __launch_bounds__(256) k_load_owned(const fs3_particle* __restrict__ in, long long n, MigArgs mig, float4* __restrict__ pos,
                                                    float4* __restrict__ vel, uint32_t* __restrict__ counts, uint32_t cap, int* __restrict__ flags) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	fs3_particle q = in[i];
	const int owner = mig_owner(mig, q.Z);
	const uint32_t k = atomicAdd(counts + owner, 1u);
	if (owner != mig.rank) return;
	if (k >= cap) { flags[3] = 1; return; }
	pos[k] = make_float4(q.X, q.Y, q.Z, __int_as_float((int)i));
	vel[k] = make_float4(q.uX, q.uY, q.uZ, 0.0f);
}
int fs3i_load_particles_owned(fs3_sim* s, const fs3_particle* dev_in, long long count) {
	s->cur = 0; s->sorted = false; s->counted = false; s->closest_valid = false;
	FS3_CUDA(cudaMemsetAsync(s->mig_ctl, 0, (2 * s->world + 1) * sizeof(uint32_t), s->stream));
	if (count > 0) {
		k_load_owned<<<cdiv(count, 256), 256, 0, s->stream>>>(dev_in, count, mig_args(s), s->pos[0] + s->poff, s->vel[0] + s->poff, s->mig_ctl,
		                                                      (uint32_t)s->cap, s->dev_flags);
		FS3_LAUNCH_CHECK();
	}
	FS3_CUDA(cudaMemcpyAsync(s->h_pin, s->mig_ctl, s->world * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	for (int r = 0; r < s->world; r++) s->np_all[r] = s->h_pin[r];
	s->np = s->np_in = std::min<long long>(s->np_all[s->rank], s->cap);
	return fs3i_check_flags(s);
}
// this rank's particles as compact AoS + creation ids (live range must be hole-free: true at every API boundary)
__global__ void k_pack_local(const float4* __restrict__ pos, const float4* __restrict__ vel, long long np, fs3_particle* __restrict__ out, int* __restrict__ ids) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i], v = vel[i];
	out[i] = fs3_particle{p.x, p.y, p.z, v.x, v.y, v.z};
	ids[i] = __float_as_int(p.w);
}
__global__ void k_load_local(const fs3_particle* __restrict__ in, const int* __restrict__ ids, long long np, float4* __restrict__ pos, float4* __restrict__ vel) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	fs3_particle q = in[i];
	pos[i] = make_float4(q.X, q.Y, q.Z, __int_as_float(ids[i]));
	vel[i] = make_float4(q.uX, q.uY, q.uZ, 0.0f);
}
int fs3i_pack_local(fs3_sim* s, fs3_particle* dev_out, int* dev_ids) {
	if (s->np_in != s->np) { s->err = "live range has holes (mid-step)"; return FS3_ERR_STATE; }
	if (s->np <= 0) return FS3_OK;
	k_pack_local<<<cdiv(s->np, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff, s->np, dev_out, dev_ids); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_load_local(fs3_sim* s, const fs3_particle* dev_in, const int* dev_ids, long long count) {
	// per-rank call (not collective): the live buffer index `cur` is shared state of the whole group and is left alone
	s->np = s->np_in = count; s->np_all[s->rank] = count; s->sorted = false; s->counted = false; s->closest_valid = false;
	if (count <= 0) return FS3_OK;
	k_load_local<<<cdiv(count, 256), 256, 0, s->stream>>>(dev_in, dev_ids, count, s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_cell_indices(fs3_sim* s, int* dev_out) {
	if (s->np_in <= 0) return FS3_OK;
	GridDims d{s->nx, s->ny, s->nz};
	k_cell_indices<<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->np_in, d, dev_out); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// ------------------------------------------------------------------------------------------------
// PIC/FLIP blend (S3D.cpp:148-180): old <- cur - (1-alpha)*old on the three face grids, then
// u_p <- (1-alpha) u_p + interp(old) per particle.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_flip_combine(const float* __restrict__ cur, float* __restrict__ old, size_t n, float om) {
	size_t q4 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	size_t n4 = n >> 2;
	if (q4 < n4) {
		float4 c = reinterpret_cast<const float4*>(cur)[q4];
		float4 o = reinterpret_cast<float4*>(old)[q4];
		o.x = __fsub_rn(c.x, __fmul_rn(om, o.x)); o.y = __fsub_rn(c.y, __fmul_rn(om, o.y));
		o.z = __fsub_rn(c.z, __fmul_rn(om, o.z)); o.w = __fsub_rn(c.w, __fmul_rn(om, o.w));
		reinterpret_cast<float4*>(old)[q4] = o;
	} else if (q4 == n4) {
		for (size_t q = n4 << 2; q < n; q++) old[q] = __fsub_rn(cur[q], __fmul_rn(om, old[q]));
	}
}
int fs3i_flip_combine(fs3_sim* s, float alpha) {
	float om = 1.0f - alpha;
	if (s->world > 1) {
		// sharded: G2P only interpolates within one cell of this slab's particles: the slab's planes and two on either side
		// (k_flip_combine handles a range that is not 16-byte aligned with its scalar tail: the range starts on a float4 boundary
		// because the first plane is rounded down to one)
		float *cur[3] = {s->U, s->V, s->W}, *old[3] = {s->oU, s->oV, s->oW};
		for (int ax = 0; ax < 3; ax++) {
			const size_t plane = (size_t)(s->nx + (ax == 0)) * (s->ny + (ax == 1));
			const int top = s->nz + (ax == 2 ? 1 : 0);
			const int a = std::max(0, s->z0 - 2), b = std::min(top, s->z1 + 2 + (ax == 2 ? 1 : 0));
			size_t q0 = plane * a, q1 = plane * b;
			q0 &= ~(size_t)3;
			k_flip_combine<<<cdiv((long long)((q1 - q0) >> 2) + 1, 256), 256, 0, s->stream>>>(cur[ax] + q0, old[ax] + q0, q1 - q0, om); FS3_LAUNCH_CHECK();
		}
		return FS3_OK;
	}
	k_flip_combine<<<cdiv((long long)(s->nU >> 2) + 1, 256), 256, 0, s->stream>>>(s->U, s->oU, s->nU, om); FS3_LAUNCH_CHECK();
	k_flip_combine<<<cdiv((long long)(s->nV >> 2) + 1, 256), 256, 0, s->stream>>>(s->V, s->oV, s->nV, om); FS3_LAUNCH_CHECK();
	k_flip_combine<<<cdiv((long long)(s->nW >> 2) + 1, 256), 256, 0, s->stream>>>(s->W, s->oW, s->nW, om); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

__global__ void __launch_bounds__(256) k_g2p(const float4* __restrict__ pos, float4* __restrict__ vel, long long np, GridDims d, float om,
                                             const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= np) return;
	float4 p = pos[i], v = vel[i];
	if (__float_as_int(p.w) < 0) return;
	float3 g = fs3_interp(U, V, W, d, __fmul_rn((float)d.nx, p.x), __fmul_rn((float)d.ny, p.y), __fmul_rn((float)d.nz, p.z));
	v.x = __fadd_rn(__fmul_rn(om, v.x), g.x);
	v.y = __fadd_rn(__fmul_rn(om, v.y), g.y);
	v.z = __fadd_rn(__fmul_rn(om, v.z), g.z);
	if (v.x > 100000.0f) v.x = 1000000.0f;  // S3D.cpp:172-175
	vel[i] = v;
}
int fs3i_g2p(fs3_sim* s, float alpha) {
	if (s->np_in <= 0) return FS3_OK;
	GridDims d{s->nx, s->ny, s->nz};
	k_g2p<<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff, s->np_in, d, 1.0f - alpha, s->oU, s->oV, s->oW);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}
