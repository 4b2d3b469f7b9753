// TransferParticlesToGrid (S3D.cpp:422-612) and ExtrapolateValues (S3D.cpp:614-778) on sm_100a.
//
// P2G is a deterministic, work-optimal GATHER over the half-cell-sorted particles.  The support of a
// trilinear-hat node is [m-1, m+1) per axis in the node's own coordinate (n*X, +0.5 on the staggered
// axis), which is exactly 4 consecutive half-cells per axis, and half-cells are stored x fastest, so one
// thread per face node walks 4 x 4 contiguous particle runs in which EVERY particle contributes.  The
// weight is evaluated as 1 - |q - m|: for q in [m-1, m+1] this is bit-identical to the reference's
// floor/alpha form (q - floor(q) and m - q are exact in fp32), and a particle whose rounded q lands on
// the support's edge gets weight 0 where the reference has no tap.  sum(w*u) and sum(w) are accumulated
// in fp32 per tap like the reference's scatter; only the order of the additions differs (run order
// instead of particle-index order).  Normalisation, the validity test and the wall faces are fused in,
// and the kernel also initialises the L1-distance grid of the extrapolation.
#include "fs3_internal.cuh"

// AX = staggered axis of this face grid (0:U, 1:V, 2:W).  Grid extents (xs,ys,zs) = cells + 1 on AX.
template <int AX>
__global__ void __launch_bounds__(128) k_p2g_gather(const float4* __restrict__ pos, const float4* __restrict__ vel,
                                                    const uint32_t* __restrict__ start, GridDims d, float cpm, float* __restrict__ grid,
                                                    int* __restrict__ cd, size_t q0, size_t q1) {
	const int xs = d.nx + (AX == 0), ys = d.ny + (AX == 1), zs = d.nz + (AX == 2);
	size_t q = q0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= q1) return;
	int i = (int)(q % xs), j = (int)((q / xs) % ys), k = (int)(q / ((size_t)xs * ys));
	// wall planes: forced to zero and valid (S3D.cpp:578-599)
	int m_ax = AX == 0 ? i : (AX == 1 ? j : k), n_ax = AX == 0 ? d.nx : (AX == 1 ? d.ny : d.nz);
	if (m_ax == 0 || m_ax == n_ax) { grid[q] = 0.f; cd[q] = 0; return; }
	// half-cell support: staggered axis [2m-2, 2m+1], other axes [2m-1, 2m+2], clipped to the grid
	const int HX = 2 * d.nx, HY = 2 * d.ny, HZ = 2 * d.nz;
	int x0 = max(2 * i - (AX == 0 ? 2 : 1), 0), x1 = min(2 * i + (AX == 0 ? 1 : 2), HX - 1);
	int y0 = max(2 * j - (AX == 1 ? 2 : 1), 0), y1 = min(2 * j + (AX == 1 ? 1 : 2), HY - 1);
	int z0 = max(2 * k - (AX == 2 ? 2 : 1), 0), z1 = min(2 * k + (AX == 2 ? 1 : 2), HZ - 1);
	const float fi = (float)i, fj = (float)j, fk = (float)k;
	const float* __restrict__ vcomp = reinterpret_cast<const float*>(vel) + AX;
	float sum = 0.f, amt = 0.f;
	for (int hz = z0; hz <= z1; hz++)
		for (int hy = y0; hy <= y1; hy++) {
			size_t row = (size_t)HX * ((size_t)hy + (size_t)HY * hz);
			uint32_t b = __ldg(start + row + x0), e = __ldg(start + row + x1 + 1);
			for (uint32_t t = b; t < e; t++) {
				float4 p = __ldg(pos + t);
				float qx = __fmul_rn(p.x, cpm), qy = __fmul_rn(p.y, cpm), qz = __fmul_rn(p.z, cpm);   // S3D.cpp:442-444
				if (AX == 0) qx = __fadd_rn(qx, 0.5f);
				if (AX == 1) qy = __fadd_rn(qy, 0.5f);
				if (AX == 2) qz = __fadd_rn(qz, 0.5f);
				float wx = __fsub_rn(1.f, fabsf(__fsub_rn(qx, fi)));
				float wy = __fsub_rn(1.f, fabsf(__fsub_rn(qy, fj)));
				float wz = __fsub_rn(1.f, fabsf(__fsub_rn(qz, fk)));
				float w = __fmul_rn(__fmul_rn(wx, wy), wz);                                           // S3D.cpp:472-474
				float u = __ldg(vcomp + 4 * (size_t)t);
				sum = __fadd_rn(sum, __fmul_rn(w, u));
				amt = __fadd_rn(amt, w);
			}
		}
	const float tiny = 1.401298464324817e-45f;  // numeric_limits<float>::denorm_min(), S3D.cpp:540
	grid[q] = __fdiv_rn(sum, __fadd_rn(tiny, amt));
	cd[q] = amt > 0.01f ? 0 : FS3_INF_CD;  // S3D.cpp:565-575
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
__global__ void __launch_bounds__(256) k_ex_scatter(const int* __restrict__ cd, size_t n, const uint32_t* __restrict__ start,
                                                    uint32_t* __restrict__ cursor, int* __restrict__ order, int nb, size_t chunk) {
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
	for (size_t q = q0 + threadIdx.x; q < q1; q += blockDim.x) {
		int c = cd[q];
		if (c > 0) { int b = min(c, nb - 1); order[cnt[b] + atomicAdd(loc + b, 1u)] = (int)q; }
	}
}

__global__ void __launch_bounds__(256) k_ex_fill(float* __restrict__ g, const int* __restrict__ cd, const int* __restrict__ order, uint32_t b,
                                                 uint32_t e, int xs, int ys, int zs) {
	uint32_t t = b + blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= e) return;
	int q = order[t];
	int my = cd[q];
	int x = q % xs, y = (q / xs) % ys, z = q / (xs * ys);
	int sy = xs, sz = xs * ys;
	float cnt = 0.f, sum = 0.f;
	// neighbour order +x,-x,+y,-y,+z,-z (S3D.cpp:743)
	if (x + 1 < xs && cd[q + 1] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q + 1]); }
	if (x - 1 >= 0 && cd[q - 1] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q - 1]); }
	if (y + 1 < ys && cd[q + sy] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q + sy]); }
	if (y - 1 >= 0 && cd[q - sy] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q - sy]); }
	if (z + 1 < zs && cd[q + sz] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q + sz]); }
	if (z - 1 >= 0 && cd[q - sz] < my) { cnt = __fadd_rn(cnt, 1.f); sum = __fadd_rn(sum, g[q - sz]); }
	g[q] = __fdiv_rn(sum, cnt);
}

int fs3i_cd_from_valid(fs3_sim* s, const unsigned char* dev_valid, size_t n) {
	k_cd_from_valid<<<cdiv((long long)n, 256), 256, 0, s->stream>>>(dev_valid, s->cdv[0], n); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// Phase 1 for scratch slot `slot` (cdv[slot] initialised to 0 / FS3_INF_CD): distance transform, bucket
// histogram + starts (copied to pinned host memory), node lists.  Asynchronous.
int fs3i_extrap_prepare(fs3_sim* s, int slot, int xs, int ys, int zs) {
	size_t n = (size_t)xs * ys * zs;
	if (n > 0x7fffffffull) { s->err = "face grid too large for 32-bit node indices"; return FS3_ERR_INVALID; }
	int* cd = s->cdv[slot];
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
	FS3_CUDA(cudaMemcpyAsync(s->h_bucket + (size_t)slot * (s->bucket_cap + 1), start, sizeof(uint32_t) * (nb + 1), cudaMemcpyDeviceToHost, s->stream));
	size_t chunk = 8192;
	k_ex_scatter<<<cdiv((long long)n, (long long)chunk), 256, 2 * nb * sizeof(uint32_t), s->stream>>>(cd, n, start, cursor, s->orderv[slot], nb, chunk);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// Phase 2: layer-by-layer fill.  Needs the bucket starts on the host (caller synchronised the stream).
int fs3i_extrap_fill(fs3_sim* s, int slot, float* grid, int xs, int ys, int zs) {
	int nb = xs + ys + zs + 1;
	const uint32_t* hb = s->h_bucket + (size_t)slot * (s->bucket_cap + 1);
	for (int dd = 1; dd < nb; dd++) {
		uint32_t b = hb[dd], e = hb[dd + 1];
		if (e > b) {
			k_ex_fill<<<cdiv((long long)(e - b), 256), 256, 0, s->stream>>>(grid, s->cdv[slot], s->orderv[slot], b, e, xs, ys, zs);
			FS3_LAUNCH_CHECK();
		}
	}
	return FS3_OK;
}

int fs3i_extrapolate(fs3_sim* s, float* grid, int xs, int ys, int zs) {
	int rc = fs3i_extrap_prepare(s, 0, xs, ys, zs);
	if (rc) return rc;
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return fs3i_extrap_fill(s, 0, grid, xs, ys, zs);
}

// Face nodes of this rank's slab: z planes [z0, z1) of every face grid, plus the top W plane on the last rank.
void fs3i_p2g_range(const fs3_sim* s, int ax, size_t* q0, size_t* q1) {
	const size_t plane = (size_t)(s->nx + (ax == 0)) * (s->ny + (ax == 1));
	*q0 = plane * s->z0;
	*q1 = plane * (s->z1 + ((ax == 2 && s->rank == s->world - 1) ? 1 : 0));
}

int fs3i_p2g_gather(fs3_sim* s) {
	GridDims d{s->nx, s->ny, s->nz};
	const float4 *pos = s->pos[s->cur], *vel = s->vel[s->cur];
	// P2G writes the "old" set; the projected field is produced from it by fs3i_project (no copy, S3D.cpp:123-134)
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_P2G], s->stream);
	size_t a, b;
	fs3i_p2g_range(s, 0, &a, &b);
	k_p2g_gather<0><<<cdiv((long long)(b - a), 128), 128, 0, s->stream>>>(pos, vel, s->cell_start, d, s->cpm, s->oU, s->cdv[0], a, b); FS3_LAUNCH_CHECK();
	fs3i_p2g_range(s, 1, &a, &b);
	k_p2g_gather<1><<<cdiv((long long)(b - a), 128), 128, 0, s->stream>>>(pos, vel, s->cell_start, d, s->cpm, s->oV, s->cdv[1], a, b); FS3_LAUNCH_CHECK();
	fs3i_p2g_range(s, 2, &a, &b);
	k_p2g_gather<2><<<cdiv((long long)(b - a), 128), 128, 0, s->stream>>>(pos, vel, s->cell_start, d, s->cpm, s->oW, s->cdv[2], a, b); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

int fs3i_p2g(fs3_sim* s) {
	int rc = fs3i_bin(s);
	if (rc) return rc;
	if ((rc = fs3i_p2g_gather(s))) return rc;
	return fs3i_extrapolate_all(s);
}

// the three face grids over the WHOLE grid (a sharded handle runs this replicated on the gathered P2G result)
int fs3i_extrapolate_all(fs3_sim* s) {
	GridDims d{s->nx, s->ny, s->nz};
	int rc;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_EXTRAPOLATE], s->stream);
	for (int ax = 0; ax < 3; ax++) {
		rc = fs3i_extrap_prepare(s, ax, d.nx + (ax == 0), d.ny + (ax == 1), d.nz + (ax == 2));
		if (rc) return rc;
	}
	FS3_CUDA(cudaStreamSynchronize(s->stream));  // one host round trip per step: layer sizes shape the fill launches
	for (int ax = 0; ax < 3; ax++) {
		rc = fs3i_extrap_fill(s, ax, ax == 0 ? s->oU : (ax == 1 ? s->oV : s->oW), d.nx + (ax == 0), d.ny + (ax == 1), d.nz + (ax == 2));
		if (rc) return rc;
	}
	return FS3_OK;
}
