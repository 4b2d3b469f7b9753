// TransferParticlesToGrid (S3D.cpp:422-612) and ExtrapolateValues (S3D.cpp:614-778) on sm_100a.
//
// P2G is a deterministic gather over the half-cell-sorted particles (no atomics): see k_p2g_units.  sum(w*u) and
// sum(w) are accumulated in fp32 per tap like the reference's scatter; only the order of the additions differs (unit order
// instead of particle-index order).  Normalisation, the validity test and the wall faces are fused in, and the kernel also
// initialises the L1-distance grid of the extrapolation.
#include "fs3_internal.cuh"
#include <algorithm>
#include <cstdlib>

#define FS3_INVALID_TAG 0x7fc0f53a   // a quiet NaN no arithmetic produces

// ------------------------------------------------------------------------------------------------
// P2G as UNIT partial sums (AX = staggered axis of the face grid, 0:U 1:V 2:W; grid extents = cells + 1 on AX).  Per face grid the particles are cut into UNITS of 2 x 2 x 2 half-cells chosen
// so that every particle of a unit has the same floor() on each axis -- the cell itself on the staggered axis (q + 0.5 in
// [c, c+1)), the interval between two cell centres on the other axes (q in [c, c+1)): all particles of a unit touch the SAME 8
// nodes, and every (particle, node) pair that is evaluated carries weight.  A thread owns one unit, walks its four particle
// runs once and keeps the 8 (sum of w*u, sum of w) pairs in registers; the 8 x 8 x 8 units of a CTA then meet in shared
// memory, where the 7 x 7 x 7 nodes that have all eight of their units in the CTA add them in a fixed order (units of the halo
// are recomputed by the neighbouring CTA: 1.49x the particle work, no global round trip, no atomics).  Deterministic, and
// independent of how the grid is cut into CTAs or z-slabs: a node is always the same 8 unit sums in the same order.
// Against the round-1 gather (a thread per 2 x 2 x 2 nodes walking their joint 6 x 6 x 6 half-cell support, 30 % of the
// evaluated pairs useful, runs of ~6 particles): no zero-weight pairs and runs of ~2 particles (less divergence), at the price
// of the 1.49x halo; 7.7 -> 6.9 ms beside the level-set sweeps at 256^3.
// Weights are 1 - (q - m) and 1 - ((m + 1) - q), bit-identical to the reference's alpha / 1 - alpha form (S3D.cpp:458-536).
#define PU_T 8                       // units per CTA edge
#define PU_N (PU_T * PU_T * PU_T)
template <int AX>
__global__ void __launch_bounds__(PU_N) k_p2g_units(const float4* __restrict__ pos, const float4* __restrict__ vel, const uint32_t* __restrict__ start,
                                                    GridDims d, float cpm, float* __restrict__ grid, int* __restrict__ cd, int kbeg, int kend, int tag_invalid) {
	__shared__ float s_sum[8][PU_N], s_amt[8][PU_N];
	const int tid = threadIdx.x, lx = tid & (PU_T - 1), ly = (tid >> 3) & (PU_T - 1), lz = tid >> 6;
	// unit coordinates = floor() of its particles in node space; the CTA's nodes are units origin+1 .. origin+7 per axis
	const int ux = (int)blockIdx.x * (PU_T - 1) - 1 + lx, uy = (int)blockIdx.y * (PU_T - 1) - 1 + ly, uz = kbeg + (int)blockIdx.z * (PU_T - 1) - 1 + lz;
	const int HX = 2 * d.nx, HY = 2 * d.ny, HZ = 2 * d.nz;
	// half-cells of the unit: staggered axis [2u, 2u+1] (the cell u), other axes [2u+1, 2u+2] (between the centres of u and u+1)
	const int hx0 = max(2 * ux + (AX == 0 ? 0 : 1), 0), hx1 = min(2 * ux + (AX == 0 ? 1 : 2), HX - 1);
	const int hy0 = max(2 * uy + (AX == 1 ? 0 : 1), 0), hy1 = min(2 * uy + (AX == 1 ? 1 : 2), HY - 1);
	const int hz0 = max(2 * uz + (AX == 2 ? 0 : 1), 0), hz1 = min(2 * uz + (AX == 2 ? 1 : 2), HZ - 1);
	const float mx = (float)ux, my = (float)uy, mz = (float)uz, mx1 = (float)(ux + 1), my1 = (float)(uy + 1), mz1 = (float)(uz + 1);
	const float* __restrict__ vcomp = reinterpret_cast<const float*>(vel) + AX;
	float sum[8], amt[8];
#pragma unroll
	for (int k = 0; k < 8; k++) sum[k] = amt[k] = 0.f;
	// only units that touch a node this launch stores (planes [kbeg, kend)): on a sharded handle the offsets beyond the ghost
	// layers of the slab are stale.  (Walking the four runs as ONE flattened loop with a one-ahead prefetch was tried: fewer warp
	// iterations, but 4.55 against 4.26 ms for the three grids alone at 256^3.)
	if (hx0 <= hx1 && uz >= kbeg - 1 && uz < kend) {
		for (int hz = hz0; hz <= hz1; hz++)
			for (int hy = hy0; hy <= hy1; hy++) {
				const size_t row = (size_t)HX * ((size_t)hy + (size_t)HY * hz);
				const uint32_t b = __ldg(start + row + hx0), e = __ldg(start + row + hx1 + 1);
				for (uint32_t t = b; t < e; t++) {
					const float4 p = __ldg(pos + t);
					const float u = __ldg(vcomp + 4 * (size_t)t);
					float qx = __fmul_rn(p.x, cpm), qy = __fmul_rn(p.y, cpm), qz = __fmul_rn(p.z, cpm);   // S3D.cpp:442-444
					if (AX == 0) qx = __fadd_rn(qx, 0.5f);
					if (AX == 1) qy = __fadd_rn(qy, 0.5f);
					if (AX == 2) qz = __fadd_rn(qz, 0.5f);
					const float wx[2] = {__fsub_rn(1.f, __fsub_rn(qx, mx)), __fsub_rn(1.f, __fsub_rn(mx1, qx))};
					const float wy[2] = {__fsub_rn(1.f, __fsub_rn(qy, my)), __fsub_rn(1.f, __fsub_rn(my1, qy))};
					const float wz[2] = {__fsub_rn(1.f, __fsub_rn(qz, mz)), __fsub_rn(1.f, __fsub_rn(mz1, qz))};
#pragma unroll
					for (int c = 0; c < 2; c++)
#pragma unroll
						for (int bb = 0; bb < 2; bb++)
#pragma unroll
							for (int a = 0; a < 2; a++) {
								const float w = __fmul_rn(__fmul_rn(wx[a], wy[bb]), wz[c]);                              // S3D.cpp:472-474
								sum[4 * c + 2 * bb + a] = __fadd_rn(sum[4 * c + 2 * bb + a], __fmul_rn(w, u));
								amt[4 * c + 2 * bb + a] = __fadd_rn(amt[4 * c + 2 * bb + a], w);
							}
				}
			}
	}
#pragma unroll
	for (int k = 0; k < 8; k++) { s_sum[k][tid] = sum[k]; s_amt[k][tid] = amt[k]; }
	__syncthreads();
	// node (i,j,k) = the upper corner of unit (i-1,..) ... the lower corner of unit (i,j,k): eight unit sums, fixed order
	if (lx == 0 || ly == 0 || lz == 0) return;
	const int xs = d.nx + (AX == 0), ys = d.ny + (AX == 1);
	const int i = ux, j = uy, k = uz;
	if (i >= xs || j >= ys || k >= kend) return;
	float S = 0.f, A = 0.f;
#pragma unroll
	for (int c = 1; c >= 0; c--)
#pragma unroll
		for (int bb = 1; bb >= 0; bb--)
#pragma unroll
			for (int a = 1; a >= 0; a--) {
				const int src = tid - a - bb * PU_T - c * PU_T * PU_T;     // unit (i-a, j-bb, k-c), whose corner (a, bb, c) is this node
				S = __fadd_rn(S, s_sum[4 * c + 2 * bb + a][src]);
				A = __fadd_rn(A, s_amt[4 * c + 2 * bb + a][src]);
			}
	const size_t q = (size_t)i + (size_t)xs * ((size_t)j + (size_t)ys * k);
	// wall planes: forced to zero and valid (S3D.cpp:578-599)
	const int m_ax = AX == 0 ? i : (AX == 1 ? j : k), n_ax = AX == 0 ? d.nx : (AX == 1 ? d.ny : d.nz);
	if (m_ax == 0 || m_ax == n_ax) { grid[q] = 0.f; cd[q] = 0; return; }
	const float tiny = 1.401298464324817e-45f;  // numeric_limits<float>::denorm_min(), S3D.cpp:540
	const bool valid = A > 0.01f;  // S3D.cpp:565-575
	grid[q] = (!valid && tag_invalid) ? __int_as_float(FS3_INVALID_TAG) : __fdiv_rn(S, __fadd_rn(tiny, A));
	cd[q] = valid ? 0 : FS3_INF_CD;
}

// ------------------------------------------------------------------------------------------------
// ExtrapolateValues.  (1) exact L1 distance transform by separable min-plus passes (the reference's six
// scans, S3D.cpp:641-699); (2) nodes with d>0 bucketed by distance; (3) layers filled from the inside
// out, each node = mean of its 6-neighbours that are strictly closer (S3D.cpp:742-769).  Nodes of one
// layer only read closer layers, so a layer is one fully parallel launch and the result does not depend
// on the order inside a bucket.
// ------------------------------------------------------------------------------------------------
__global__ void k_cd_from_valid(const unsigned char* __restrict__ valid, int* __restrict__ cd, size_t n) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < n) cd[q] = valid[q] ? 0 : FS3_INF_CD;
}

// x passes: one warp per row, prefix/suffix min of (cd -/+ x) with shuffles, coalesced.
__global__ void __launch_bounds__(256) k_dt_x(int* __restrict__ cd, int xs, size_t rows) {
	size_t row = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
	int lane = threadIdx.x & 31;
	if (row >= rows) return;
	int* r = cd + row * (size_t)xs;
	const int BIG = 0x7f000000;
	int carry = BIG;
	for (int c0 = 0; c0 < xs; c0 += 32) {
		int x = c0 + lane;
		int v = x < xs ? r[x] - x : BIG;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v = min(v, t); }
		v = min(v, carry);
		carry = __shfl_sync(0xffffffffu, v, 31);
		if (x < xs) r[x] = v + x;
	}
	carry = BIG;
	for (int c0 = ((xs - 1) / 32) * 32; c0 >= 0; c0 -= 32) {
		int x = c0 + lane;
		int v = x < xs ? r[x] + x : BIG;
#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { int t = __shfl_down_sync(0xffffffffu, v, o); if (lane + o < 32) v = min(v, t); }
		v = min(v, carry);
		carry = __shfl_sync(0xffffffffu, v, 0);
		if (x < xs) r[x] = v - x;
	}
}

// y or z passes: one thread per line, forward then backward march; loads batched 8 deep for ILP.
__global__ void __launch_bounds__(128) k_dt_line(int* __restrict__ cd, int inner, size_t outer_stride, size_t nlines, int len, size_t stride) {
	size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= nlines) return;
	int* base = cd + (t % inner) + (t / inner) * outer_stride;
	int prev = FS3_INF_CD;
	for (int y0 = 0; y0 < len; y0 += 8) {
		int v[8];
#pragma unroll
		for (int u = 0; u < 8; u++) v[u] = (y0 + u < len) ? base[(size_t)(y0 + u) * stride] : FS3_INF_CD;
#pragma unroll
		for (int u = 0; u < 8; u++) { prev = min(v[u], prev + 1); if (y0 + u < len) base[(size_t)(y0 + u) * stride] = prev; }
	}
	prev = FS3_INF_CD;
	for (int y0 = len - 1; y0 >= 0; y0 -= 8) {
		int v[8];
#pragma unroll
		for (int u = 0; u < 8; u++) v[u] = (y0 - u >= 0) ? base[(size_t)(y0 - u) * stride] : FS3_INF_CD;
#pragma unroll
		for (int u = 0; u < 8; u++) { prev = min(v[u], prev + 1); if (y0 - u >= 0) base[(size_t)(y0 - u) * stride] = prev; }
	}
}

// histogram of distances (cd>0 only), privatised in shared memory
__global__ void __launch_bounds__(256) k_ex_hist(const int* __restrict__ cd, size_t n, uint32_t* __restrict__ hist, int nb) {
	extern __shared__ uint32_t sh[];
	for (int b = threadIdx.x; b < nb; b += blockDim.x) sh[b] = 0;
	__syncthreads();
	for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (size_t)gridDim.x * blockDim.x) {
		int c = cd[q];
		if (c > 0) atomicAdd(sh + min(c, nb - 1), 1u);
	}
	__syncthreads();
	for (int b = threadIdx.x; b < nb; b += blockDim.x) if (sh[b]) atomicAdd(hist + b, sh[b]);
}
// exclusive scan of the histogram (one block); start[] gets nb+1 entries, cursor[] zeroed
__global__ void k_ex_scan(const uint32_t* __restrict__ hist, uint32_t* __restrict__ start, uint32_t* __restrict__ cursor, int nb) {
	if (threadIdx.x == 0) {
		uint32_t acc = 0;
		for (int b = 0; b < nb; b++) { start[b] = acc; acc += hist[b]; cursor[b] = 0; }
		start[nb] = acc;
	}
}
// Node lists bucketed by distance.  The scatter also records, per list entry, which of the node's 6 neighbours
// (+x,-x,+y,-y,+z,-z, S3D.cpp:743) are strictly closer: it walks the grid in memory order, so the neighbour distances
// come from cache here, and the fill -- whose layers are one dependent step each -- is left with the value loads only.
__global__ void __launch_bounds__(256) k_ex_scatter(const int* __restrict__ cd, size_t n, const uint32_t* __restrict__ start,
                                                    uint32_t* __restrict__ cursor, int* __restrict__ order, unsigned char* __restrict__ exmask,
                                                    int nb, size_t chunk, int xs, int ys, int zs) {
	extern __shared__ uint32_t sh[];  // [0,nb): block counts then block base; [nb,2nb): local cursor
	uint32_t* cnt = sh;
	uint32_t* loc = sh + nb;
	for (int b = threadIdx.x; b < 2 * nb; b += blockDim.x) sh[b] = 0;
	__syncthreads();
	size_t q0 = (size_t)blockIdx.x * chunk, q1 = min(n, q0 + chunk);
	for (size_t q = q0 + threadIdx.x; q < q1; q += blockDim.x) { int c = cd[q]; if (c > 0) atomicAdd(cnt + min(c, nb - 1), 1u); }
	__syncthreads();
	for (int b = threadIdx.x; b < nb; b += blockDim.x) if (cnt[b]) cnt[b] = start[b] + atomicAdd(cursor + b, cnt[b]);
	__syncthreads();
	const size_t sy = (size_t)xs, sz = (size_t)xs * ys;
	for (size_t q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
		const int my = cd[q];
		if (my > 0) {
			const int x = (int)(q % xs), y = (int)((q / xs) % ys), z = (int)(q / sz);
			int m = 0;
			if (x + 1 < xs && __ldg(cd + q + 1) < my) m |= 1;
			if (x - 1 >= 0 && __ldg(cd + q - 1) < my) m |= 2;
			if (y + 1 < ys && __ldg(cd + q + sy) < my) m |= 4;
			if (y - 1 >= 0 && __ldg(cd + q - sy) < my) m |= 8;
			if (z + 1 < zs && __ldg(cd + q + sz) < my) m |= 16;
			if (z - 1 >= 0 && __ldg(cd + q - sz) < my) m |= 32;
			const int b = min(my, nb - 1);
			const uint32_t slot = cnt[b] + atomicAdd(loc + b, 1u);
			order[slot] = (int)q;
			exmask[slot] = (unsigned char)m;
		}
	}
}

// All layers of up to three face grids in ONE persistent launch.  Layer d only reads layer d-1 (a node's closer
// neighbours are exactly one step closer in L1 distance), so the layers are separated by a grid-wide barrier (arrive
// counter in L2, one CTA per SM, all co-resident) instead of a kernel boundary, and the bucket starts never leave the
// device.  Values of the previous layer are read with ld.cg (other SMs wrote them); cd[] is constant during the fill.
struct ExGrid {
	float* g;
	const int* order;
	const unsigned char* mask;   // closer-neighbour bits of the list entries
	const uint32_t* start;   // nb+1 bucket starts
	int xs, ys, zs, nb;
};
struct ExGrids { ExGrid a[3]; int n; };
#define EX_THREADS 512

__device__ __forceinline__ void ex_fill_node(const ExGrid& G, int q, int m) {
	float* __restrict__ g = G.g;
	const int sy = G.xs, sz = G.xs * G.ys;
	// all loads first (one L2 round trip), then the reference's sequence of additions
	const float v0 = (m & 1) ? __ldcg(g + q + 1) : 0.f, v1 = (m & 2) ? __ldcg(g + q - 1) : 0.f;
	const float v2 = (m & 4) ? __ldcg(g + q + sy) : 0.f, v3 = (m & 8) ? __ldcg(g + q - sy) : 0.f;
	const float v4 = (m & 16) ? __ldcg(g + q + sz) : 0.f, v5 = (m & 32) ? __ldcg(g + q - sz) : 0.f;
	float cnt = 0.f, sum = 0.f;
	if (m & 1) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v0); }
	if (m & 2) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v1); }
	if (m & 4) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v2); }
	if (m & 8) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v3); }
	if (m & 16) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v4); }
	if (m & 32) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, v5); }
	__stcg(g + q, __fdiv_rn(sum, cnt));
}

__global__ void __launch_bounds__(EX_THREADS, 3) k_ex_fill_layers(ExGrids P, uint32_t* __restrict__ arrive, int* __restrict__ flags, int max_layers) {
	const uint32_t tid = blockIdx.x * EX_THREADS + threadIdx.x, nthr = gridDim.x * EX_THREADS;
	__shared__ int s_abort;
	if (threadIdx.x == 0) s_abort = 0;
	uint32_t tot[3], b[3], e[3];
	int pq[3], pm[3];   // this thread's first node of the coming layer and its closer-mask, fetched BEFORE the barrier
#pragma unroll
	for (int k = 0; k < 3; k++) {
		tot[k] = k < P.n ? __ldg(P.a[k].start + P.a[k].nb) : 0u;
		b[k] = k < P.n ? __ldg(P.a[k].start + 1) : 0u;
		e[k] = (k < P.n && 1 < P.a[k].nb) ? __ldg(P.a[k].start + 2) : b[k];
		pq[k] = -1; pm[k] = 0;
		if (k < P.n && b[k] + tid < e[k]) { pq[k] = __ldg(P.a[k].order + b[k] + tid); pm[k] = __ldg(P.a[k].mask + b[k] + tid); }
	}
	for (int d = 1;; d++) {
		bool more = false;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			if (k >= P.n) continue;
			more |= e[k] < tot[k] && d < max_layers;
			// the list entry of the next iteration is fetched while this one's values are in flight
			int q = pq[k], m = pm[k];
			for (uint32_t t = b[k] + tid; t < e[k]; t += nthr) {
				int nq = -1, nm = 0;
				if (t + nthr < e[k]) { nq = __ldg(P.a[k].order + t + nthr); nm = __ldg(P.a[k].mask + t + nthr); }
				ex_fill_node(P.a[k], q, m);
				q = nq; m = nm;
			}
		}
		if (!more) break;   // grid-uniform: every CTA reads the same starts
#pragma unroll
		for (int k = 0; k < 3; k++) {
			b[k] = e[k];
			e[k] = (k < P.n && d + 1 < P.a[k].nb) ? __ldg(P.a[k].start + d + 2) : b[k];
			pq[k] = -1;
			if (k < P.n && b[k] + tid < e[k]) { pq[k] = __ldg(P.a[k].order + b[k] + tid); pm[k] = __ldg(P.a[k].mask + b[k] + tid); }
		}
		// ---- grid barrier d ----
		__syncthreads();
		if (threadIdx.x == 0) {
			__threadfence();
			atomicAdd(arrive, 1u);
			const uint32_t need = (uint32_t)d * gridDim.x;
			uint32_t v;
			// The launch is cooperative, so every CTA is resident and the spin ends; the clock bound only turns a broken
			// assumption (or a CTA that gave up) into an error report instead of a hang.
			const long long t0 = clock64();
			for (unsigned it = 1;; it++) {
				asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(arrive) : "memory");
				if (v >= need) break;
				if ((it & 4095u) == 0 && (*reinterpret_cast<volatile int*>(flags + 1) || clock64() - t0 > 8000000000LL)) { flags[1] = 1; s_abort = 1; break; }
			}
		}
		__syncthreads();
		if (s_abort) return;
	}
}

int fs3i_cd_from_valid(fs3_sim* s, const unsigned char* dev_valid, size_t n) {
	k_cd_from_valid<<<cdiv((long long)n, 256), 256, 0, s->stream>>>(dev_valid, s->cdv[0], n); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// Phase 1 for scratch slot `slot` (cdv[slot] initialised to 0 / FS3_INF_CD): distance transform, bucket
// histogram + starts (copied to pinned host memory), node lists.  Asynchronous.
int fs3i_extrap_prepare(fs3_sim* s, int slot, int xs, int ys, int zs, size_t off) {
	size_t n = (size_t)xs * ys * zs;
	if (n > 0x7fffffffull) { s->err = "face grid too large for 32-bit node indices"; return FS3_ERR_INVALID; }
	int* cd = s->cdv[slot] + off;   // off: first node of a z window of the grid (sharded handles), node indices are relative to it
	size_t rows = (size_t)ys * zs;
	k_dt_x<<<cdiv((long long)rows * 32, 256), 256, 0, s->stream>>>(cd, xs, rows); FS3_LAUNCH_CHECK();
	k_dt_line<<<cdiv((long long)xs * zs, 128), 128, 0, s->stream>>>(cd, xs, (size_t)xs * ys, (size_t)xs * zs, ys, (size_t)xs); FS3_LAUNCH_CHECK();
	k_dt_line<<<cdiv((long long)xs * ys, 128), 128, 0, s->stream>>>(cd, xs * ys, 0, (size_t)xs * ys, zs, (size_t)xs * ys); FS3_LAUNCH_CHECK();
	int nb = xs + ys + zs + 1;
	if (nb > s->bucket_cap) { s->err = "bucket scratch too small"; return FS3_ERR_STATE; }
	uint32_t* base = s->bucket + (size_t)slot * (3 * (size_t)s->bucket_cap + 2);
	uint32_t *hist = base, *start = base + s->bucket_cap, *cursor = base + 2 * (size_t)s->bucket_cap + 1;
	FS3_CUDA(cudaMemsetAsync(hist, 0, sizeof(uint32_t) * nb, s->stream));
	int hb = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
	k_ex_hist<<<hb, 256, nb * sizeof(uint32_t), s->stream>>>(cd, n, hist, nb); FS3_LAUNCH_CHECK();
	k_ex_scan<<<1, 32, 0, s->stream>>>(hist, start, cursor, nb); FS3_LAUNCH_CHECK();
	size_t chunk = 8192;
	k_ex_scatter<<<cdiv((long long)n, (long long)chunk), 256, 2 * nb * sizeof(uint32_t), s->stream>>>(cd, n, start, cursor, s->orderv[slot], s->exmask[slot], nb, chunk, xs, ys, zs);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// Phase 2: the layered fill of `count` face grids (slots 0..count-1) in one persistent launch.
static int extrap_fill_layers(fs3_sim* s, int count, float* const* grids, const int (*dims)[3], int max_layers = 0x7fffffff) {
	ExGrids P;
	P.n = count;
	for (int k = 0; k < count; k++) {
		const uint32_t* base = s->bucket + (size_t)k * (3 * (size_t)s->bucket_cap + 2);
		P.a[k] = ExGrid{grids[k], s->orderv[k], s->exmask[k], base + s->bucket_cap, dims[k][0], dims[k][1], dims[k][2], dims[k][0] + dims[k][1] + dims[k][2] + 1};
	}
	for (int k = count; k < 3; k++) P.a[k] = P.a[0];
	FS3_CUDA(cudaMemsetAsync(s->ex_arrive, 0, sizeof(uint32_t), s->stream));
	// The grid barrier needs every CTA resident: a COOPERATIVE launch makes that all-or-nothing (the driver only starts the grid
	// when all of it fits, also beside other handles or contexts on the device).  One CTA per SM, 512 threads x <= 40 registers:
	// fits beside the resident CTAs of the level-set sweep running on stream2.
	void* args[] = {(void*)&P, (void*)&s->ex_arrive, (void*)&s->dev_flags, (void*)&max_layers};
	FS3_CUDA(cudaLaunchCooperativeKernel((const void*)k_ex_fill_layers, dim3(s->sm_count), dim3(EX_THREADS), args, 0, s->stream));
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

int fs3i_extrap_fill(fs3_sim* s, int slot, float* grid, int xs, int ys, int zs) {
	if (slot != 0) { s->err = "single-grid fill uses slot 0"; return FS3_ERR_INVALID; }
	float* g[1] = {grid};
	const int dims[1][3] = {{xs, ys, zs}};
	return extrap_fill_layers(s, 1, g, dims);
}

// ------------------------------------------------------------------------------------------------
// fs3_params.extrapolation_mode = 1: the reference's GPU formulation of the extrapolation,
// FX/gpExtrapolateParticleVelocities.hlsl:16-42 -- an invalid node becomes the mean of its valid 6-neighbours
// (sum in the order +x -x +y -y +z -z, divided by their number), or 0 when it has none; ONE layer, no distance
// transform, no buckets, no barrier per layer.  Out-of-range neighbours are texture reads that return 0 in D3D11,
// which is finite, so they count as valid zeros (kept).  The shader updates the grid in place while neighbours read
// it; reading taken here: only nodes valid BEFORE the pass are neighbours (they are never written, so the in-place
// update below is race-free).  cd = 0 marks a valid node, as everywhere in this file.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_ex_one_layer(float* __restrict__ grid, const int* __restrict__ cd, int xs, int ys, int zs) {
	const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= (size_t)xs * ys * zs || cd[q] == 0) return;
	const int x = (int)(q % xs), y = (int)((q / xs) % ys), z = (int)(q / ((size_t)xs * ys));
	const size_t plane = (size_t)xs * ys;
	float sum = 0.0f, num = 0.0f;
#define FS3_TAP(inside, off)                                                              \
	if (!(inside)) num = __fadd_rn(num, 1.0f);                                            \
	else if (cd[off] == 0) { sum = __fadd_rn(sum, grid[off]); num = __fadd_rn(num, 1.0f); }
	FS3_TAP(x + 1 < xs, q + 1)
	FS3_TAP(x > 0, q - 1)
	FS3_TAP(y + 1 < ys, q + xs)
	FS3_TAP(y > 0, q - xs)
	FS3_TAP(z + 1 < zs, q + plane)
	FS3_TAP(z > 0, q - plane)
#undef FS3_TAP
	grid[q] = num != 0.0f ? __fdiv_rn(sum, num) : 0.0f;
}
static int extrap_one_layer(fs3_sim* s, float* grid, const int* cd, int xs, int ys, int zs) {
	k_ex_one_layer<<<cdiv((long long)xs * ys * zs, 256), 256, 0, s->stream>>>(grid, cd, xs, ys, zs); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

int fs3i_extrapolate(fs3_sim* s, float* grid, int xs, int ys, int zs) {
	if (s->prm.extrapolation_mode == 1) return extrap_one_layer(s, grid, s->cdv[0], xs, ys, zs);
	int rc = fs3i_extrap_prepare(s, 0, xs, ys, zs);
	if (rc) return rc;
	return fs3i_extrap_fill(s, 0, grid, xs, ys, zs);
}

// Face nodes of this rank's slab: z planes [z0, z1) of every face grid, plus the top W plane on the last rank.
void fs3i_p2g_range(const fs3_sim* s, int ax, size_t* q0, size_t* q1) {
	const size_t plane = (size_t)(s->nx + (ax == 0)) * (s->ny + (ax == 1));
	*q0 = plane * s->z0;
	*q1 = plane * (s->z1 + ((ax == 2 && s->rank == s->world - 1) ? 1 : 0));
}

template <int AX> static int p2g_launch(fs3_sim* s, float* grid, int* cd) {
	GridDims d{s->nx, s->ny, s->nz};
	const int xs = s->nx + (AX == 0), ys = s->ny + (AX == 1);
	const int kbeg = s->z0, kend = s->z1 + ((AX == 2 && s->rank == s->world - 1) ? 1 : 0);
	dim3 g(cdiv(xs, PU_T - 1), cdiv(ys, PU_T - 1), cdiv(kend - kbeg, PU_T - 1));
	k_p2g_units<AX><<<g, PU_N, 0, s->stream>>>(s->pos[s->cur], s->vel[s->cur], s->cell_start, d, s->cpm, grid, cd, kbeg, kend, s->world > 1 ? 1 : 0);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// one face grid (0: U, 1: V, 2: W) of this rank's slab
int fs3i_p2g_gather_axis(fs3_sim* s, int ax) {
	// P2G writes the "old" set; the projected field is produced from it by fs3i_project (no copy, S3D.cpp:123-134)
	if (ax == 0 && s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_P2G], s->stream);
	return ax == 0 ? p2g_launch<0>(s, s->oU, s->cdv[0]) : ax == 1 ? p2g_launch<1>(s, s->oV, s->cdv[1]) : p2g_launch<2>(s, s->oW, s->cdv[2]);
}
int fs3i_p2g_gather(fs3_sim* s) {
	int rc;
	if ((rc = fs3i_p2g_gather_axis(s, 0)) || (rc = fs3i_p2g_gather_axis(s, 1)) || (rc = fs3i_p2g_gather_axis(s, 2))) return rc;
	return FS3_OK;
}

// Sharded: validity (cd = 0 / inf) of the face nodes OUTSIDE this rank's slab from the tags in the gathered grids.
__global__ void __launch_bounds__(256) k_cd_from_tags(const float* __restrict__ grid, int* __restrict__ cd, size_t n, size_t own0, size_t own1) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= n - (own1 - own0)) return;
	if (q >= own0) q += own1 - own0;
	cd[q] = __float_as_int(grid[q]) == FS3_INVALID_TAG ? FS3_INF_CD : 0;
}
int fs3i_cd_from_tags(fs3_sim* s) {
	float* g[3] = {s->oU, s->oV, s->oW};
	for (int ax = 0; ax < 3; ax++) {
		size_t a, b;
		fs3i_p2g_range(s, ax, &a, &b);
		// only the planes of the extrapolation window that the neighbours pushed
		int w0, w1;
		fs3i_ex_window(s, ax, FS3_EX_DEPTH, &w0, &w1);
		const size_t plane = (size_t)(s->nx + (ax == 0)) * (s->ny + (ax == 1)), base = plane * w0, n = plane * (w1 - w0);
		const size_t rest = n - (b - a);
		if (!rest) continue;
		k_cd_from_tags<<<cdiv((long long)rest, 256), 256, 0, s->stream>>>(g[ax] + base, s->cdv[ax] + base, n, a - base, b - base); FS3_LAUNCH_CHECK();
	}
	return FS3_OK;
}

int fs3i_p2g(fs3_sim* s) {
	int rc = fs3i_bin(s);
	if (rc) return rc;
	if ((rc = fs3i_p2g_gather(s))) return rc;
	return fs3i_extrapolate_all(s);
}

// ---- sharded handles: extrapolation on a z WINDOW of the slab, FS3_EX_DEPTH layers deep -------------------------------
// A node of layer d (L1 distance d to the nearest valid node) only reads nodes of layer d-1, and every layer-d node of the slab
// lies within d planes of a valid node: with FS3_EX_DEPTH planes of the neighbours' P2G result on either side (values and
// validity, pushed by the owners) the distance transform and the first FS3_EX_DEPTH layers come out bit-identical to the
// whole-grid computation for every node of the slab -- and for the nodes k planes outside it up to layer FS3_EX_DEPTH - k.
// Deeper layers are never consumed: the divergence and the pressure gradient read faces of fluid cells (phi < 0: a particle
// within one cell of the centre, hence a valid node within L1 distance 4 of the face), air-air faces come out of the gradient as
// 0 whatever was extrapolated (S3D.cpp:1024-1037), and G2P interpolates old' = new - (1-alpha) old within one cell of a
// particle (L1 distance <= 4 again; the combine runs two planes into the halo, where layers up to FS3_EX_DEPTH - 2 are exact).
// So the slabs need no per-layer exchange and no dependency along z; nodes deeper than FS3_EX_DEPTH are set to 0.
void fs3i_ex_window(const fs3_sim* s, int ax, int depth, int* w0, int* w1) {
	const int top = s->nz + (ax == 2 ? 1 : 0);
	const int own1 = s->z1 + ((ax == 2 && s->rank == s->world - 1) ? 1 : 0);
	*w0 = std::max(0, s->z0 - depth);
	*w1 = std::min(top, own1 + depth);
}
__global__ void __launch_bounds__(256) k_ex_clear_deep(float* __restrict__ grid, const int* __restrict__ cd, size_t q0, size_t q1, int depth) {
	const size_t q = q0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < q1 && cd[q] > depth) grid[q] = 0.0f;
}
static int extrapolate_window(fs3_sim* s) {
	int rc;
	float* g[3] = {s->oU, s->oV, s->oW};
	float* gw[3];
	int dims[3][3];
	size_t off[3];
	for (int ax = 0; ax < 3; ax++) {
		const int xs = s->nx + (ax == 0), ys = s->ny + (ax == 1);
		int w0, w1;
		fs3i_ex_window(s, ax, FS3_EX_DEPTH, &w0, &w1);
		off[ax] = (size_t)xs * ys * w0;
		gw[ax] = g[ax] + off[ax];
		dims[ax][0] = xs; dims[ax][1] = ys; dims[ax][2] = w1 - w0;
		if (s->prm.extrapolation_mode == 1) {
			if ((rc = extrap_one_layer(s, gw[ax], s->cdv[ax] + off[ax], xs, ys, w1 - w0))) return rc;
		} else if ((rc = fs3i_extrap_prepare(s, ax, xs, ys, w1 - w0, off[ax]))) return rc;
	}
	if (s->prm.extrapolation_mode == 1) return FS3_OK;
	if ((rc = extrap_fill_layers(s, 3, gw, dims, FS3_EX_DEPTH))) return rc;
	for (int ax = 0; ax < 3; ax++) {
		size_t q0, q1;
		fs3i_p2g_range(s, ax, &q0, &q1);
		// (the halo planes keep whatever the fill left: exact up to their own depth, never read beyond it)
		const size_t plane = (size_t)dims[ax][0] * dims[ax][1];
		q0 = off[ax]; q1 = off[ax] + plane * dims[ax][2];
		k_ex_clear_deep<<<cdiv((long long)(q1 - q0), 256), 256, 0, s->stream>>>(g[ax], s->cdv[ax], q0, q1, FS3_EX_DEPTH); FS3_LAUNCH_CHECK();
	}
	return FS3_OK;
}

// the three face grids over the WHOLE grid (single GPU), or over the slab's window (sharded)
int fs3i_extrapolate_all(fs3_sim* s) {
	GridDims d{s->nx, s->ny, s->nz};
	int rc;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_EXTRAPOLATE], s->stream);
	if (s->world > 1) return extrapolate_window(s);
	if (s->prm.extrapolation_mode == 1) {
		float* gg[3] = {s->oU, s->oV, s->oW};
		for (int ax = 0; ax < 3; ax++)
			if ((rc = extrap_one_layer(s, gg[ax], s->cdv[ax], d.nx + (ax == 0), d.ny + (ax == 1), d.nz + (ax == 2)))) return rc;
		return FS3_OK;
	}
	for (int ax = 0; ax < 3; ax++) {
		rc = fs3i_extrap_prepare(s, ax, d.nx + (ax == 0), d.ny + (ax == 1), d.nz + (ax == 2));
		if (rc) return rc;
	}
	float* g[3] = {s->oU, s->oV, s->oW};
	const int dims[3][3] = {{d.nx + 1, d.ny, d.nz}, {d.nx, d.ny + 1, d.nz}, {d.nx, d.ny, d.nz + 1}};
	if ((rc = extrap_fill_layers(s, 3, g, dims))) return rc;
	return FS3_OK;
}
