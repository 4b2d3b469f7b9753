// C ABI, part 2: single stages of Simulate and the parity / measurement hooks (include/fs3.h).
#include "fs3_internal.cuh"
#include <cmath>
#include <vector>

#define ENTER(s)                                                       \
	if (!(s)) return FS3_ERR_INVALID;                                  \
	do {                                                               \
		cudaError_t e0_ = cudaSetDevice((s)->device);                  \
		if (e0_ != cudaSuccess) { (s)->err = cudaGetErrorString(e0_); return FS3_ERR_CUDA; } \
	} while (0)

static int fail(fs3_sim* s, int code, const char* msg) { s->err = msg; return code; }
// stage-by-stage hooks drive one GPU; a sharded handle only steps as a whole (its stages are collective phases)
#define SINGLE(s) if ((s)->world > 1) return fail(s, FS3_ERR_STATE, "this hook is not available on a sharded handle")

static float flip_alpha(const fs3_sim* s, float dt) {
	float a = 6 * dt * s->prm.nu * s->cpm * s->cpm;   // S3D.cpp:114
	return a < 0.0f ? 0.0f : (a > 1.0f ? 1.0f : a);
}

__global__ void k_fill_f32(float* p, size_t n, float v) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < n) p[q] = v;
}
__global__ void k_cell_counts(const uint32_t* __restrict__ start, GridDims d, int* __restrict__ out) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= (size_t)d.nx * d.ny * d.nz) return;
	int x = (int)(q % d.nx), y = (int)((q / d.nx) % d.ny), z = (int)(q / ((size_t)d.nx * d.ny));
	int n = 0;
	FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) n += (int)(e - b);
	out[q] = n;
}

// ---- PrintDivergence (S3D.cpp:1095-1138) and the state fingerprints ------------------------------------------------
// Block partials in a fixed tree order, then one block adds the partials in index order: the result does not depend on
// scheduling (no floating-point atomics).
__global__ void __launch_bounds__(256) k_divergence(const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W,
                                                    const float* __restrict__ phi, GridDims d, size_t c0, size_t c1, double* __restrict__ part) {
	const size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	double sq = 0.0, mx = 0.0;
	if (c < c1 && phi[c] < 0.0f) {
		const int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
		const size_t iu = (size_t)x + (size_t)(d.nx + 1) * ((size_t)y + (size_t)d.ny * z);
		const size_t iv = (size_t)x + (size_t)d.nx * ((size_t)y + (size_t)(d.ny + 1) * z);
		// velUp - velDown + velRight - velLeft + velFwds - velBkwds, left to right in fp32 (S3D.cpp:1118)
		float dv = __fsub_rn(V[iv + d.nx], V[iv]);
		dv = __fadd_rn(dv, U[iu + 1]); dv = __fsub_rn(dv, U[iu]);
		dv = __fadd_rn(dv, W[c + (size_t)d.nx * d.ny]); dv = __fsub_rn(dv, W[c]);
		mx = dv > 0.0f ? (double)dv : 0.0;
		sq = (double)__fmul_rn(dv, dv);
	}
	__shared__ double s_sq[256], s_mx[256];
	s_sq[threadIdx.x] = sq; s_mx[threadIdx.x] = mx;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) {
		if (threadIdx.x < o) { s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; s_mx[threadIdx.x] = fmax(s_mx[threadIdx.x], s_mx[threadIdx.x + o]); }
		__syncthreads();
	}
	if (threadIdx.x == 0) { part[2 * (size_t)blockIdx.x] = s_sq[0]; part[2 * (size_t)blockIdx.x + 1] = s_mx[0]; }
}
// out[0] = max of the partial maxima, out[1] = sum of the partial sums, both in a fixed order
__global__ void __launch_bounds__(256) k_partials_final(const double* __restrict__ part, size_t nblocks, double* __restrict__ out) {
	__shared__ double s_sq[256], s_mx[256];
	double sq = 0.0, mx = 0.0;
	for (size_t b = threadIdx.x; b < nblocks; b += 256) { sq += part[2 * b]; mx = fmax(mx, part[2 * b + 1]); }
	s_sq[threadIdx.x] = sq; s_mx[threadIdx.x] = mx;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) {
		if (threadIdx.x < o) { s_sq[threadIdx.x] += s_sq[threadIdx.x + o]; s_mx[threadIdx.x] = fmax(s_mx[threadIdx.x], s_mx[threadIdx.x + o]); }
		__syncthreads();
	}
	if (threadIdx.x == 0) { out[0] = s_mx[0]; out[1] = s_sq[0]; }
}

__device__ __forceinline__ unsigned long long fs3_mix64(unsigned long long x) {   // splitmix64 finaliser
	x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31;
	return x;
}
// sum over the block (256 threads), then one atomic per block: integer addition modulo 2^64 is order-independent
__device__ __forceinline__ void fs3_warp_add_u64(unsigned long long v, unsigned long long* dst) {
	__shared__ unsigned long long s_w[8];
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
	__syncthreads();   // (s_w may still be read from the previous call)
	if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
	__syncthreads();
	if (threadIdx.x == 0) {
		unsigned long long t = 0;
		for (int k = 0; k < 8; k++) t += s_w[k];
		if (t) atomicAdd(dst, t);
	}
}
__global__ void __launch_bounds__(256) k_check_particles(const float4* __restrict__ pos, const float4* __restrict__ vel, long long np, GridDims d,
                                                         unsigned long long* __restrict__ out) {
	const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long hp = 0, hc = 0, cnt = 0;
	if (i < np) {
		const float4 p = pos[i], v = vel[i];
		const int id = __float_as_int(p.w);
		if (id >= 0) {
			const int cx = fs3_cell_axis(p.x, (float)d.nx), cy = fs3_cell_axis(p.y, (float)d.ny), cz = fs3_cell_axis(p.z, (float)d.nz);
			const unsigned cell = (unsigned)(cx + d.nx * (cy + d.ny * cz));
			hc = fs3_mix64(((unsigned long long)(unsigned)id << 32) | cell);
			unsigned long long h = hc;
			h = fs3_mix64(h ^ (((unsigned long long)__float_as_uint(p.x) << 32) | __float_as_uint(p.y)));
			h = fs3_mix64(h ^ (((unsigned long long)__float_as_uint(p.z) << 32) | __float_as_uint(v.x)));
			h = fs3_mix64(h ^ (((unsigned long long)__float_as_uint(v.y) << 32) | __float_as_uint(v.z)));
			hp = h; cnt = 1;
		}
	}
	fs3_warp_add_u64(hp, out + 0); fs3_warp_add_u64(hc, out + 1); fs3_warp_add_u64(cnt, out + 2);
}
__global__ void __launch_bounds__(256) k_check_grid(const float* __restrict__ g, size_t q0, size_t q1, unsigned long long* __restrict__ hash, double* __restrict__ part) {
	const size_t q = q0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long h = 0;
	double v = 0.0;
	if (q < q1) { const float f = g[q]; h = fs3_mix64(((unsigned long long)q << 32) ^ (unsigned long long)__float_as_uint(f) ^ 0x5851f42d4c957f2dull); v = (double)f; }
	fs3_warp_add_u64(h, hash);
	__shared__ double s_v[256];
	s_v[threadIdx.x] = v;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) s_v[threadIdx.x] += s_v[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0) { part[2 * (size_t)blockIdx.x] = s_v[0]; part[2 * (size_t)blockIdx.x + 1] = 0.0; }
}

extern "C" {

int fs3_get_divergence(fs3_sim* s, double* maxdiv, double* l2) {
	ENTER(s);
	if (!maxdiv || !l2) return FS3_ERR_INVALID;
	const size_t plane = (size_t)s->nx * s->ny, c0 = plane * s->z0, c1 = plane * s->z1;
	const size_t nb = (c1 - c0 + 255) / 256;
	int rc = fs3i_staging(s, (2 * nb + 2) * sizeof(double));
	if (rc) return rc;
	double* part = (double*)s->stage_ptr;
	k_divergence<<<(unsigned)nb, 256, 0, s->stream>>>(s->U, s->V, s->W, s->phi, GridDims{s->nx, s->ny, s->nz}, c0, c1, part + 2); FS3_LAUNCH_CHECK();
	k_partials_final<<<1, 256, 0, s->stream>>>(part + 2, nb, part); FS3_LAUNCH_CHECK();
	double h[2];
	FS3_CUDA(cudaMemcpyAsync(h, part, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	*maxdiv = h[0];
	*l2 = std::sqrt(h[1]);
	return FS3_OK;
}

int fs3_checksum(fs3_sim* s, fs3_check* out) {
	ENTER(s);
	if (!out) return FS3_ERR_INVALID;
	size_t nmax = s->nW > s->nU ? s->nW : s->nU;
	if (s->nV > nmax) nmax = s->nV;
	const size_t nbmax = (nmax + 255) / 256;
	int rc = fs3i_staging(s, 64 + (2 * nbmax + 2) * sizeof(double));
	if (rc) return rc;
	unsigned long long* acc = (unsigned long long*)s->stage_ptr;            // [0..2] particles, [3..6] grid hashes
	double* part = (double*)((char*)s->stage_ptr + 64);
	FS3_CUDA(cudaMemsetAsync(acc, 0, 64, s->stream));
	const GridDims d{s->nx, s->ny, s->nz};
	if (s->np_in > 0) {
		k_check_particles<<<cdiv(s->np_in, 256), 256, 0, s->stream>>>(s->pos[s->cur] + s->poff, s->vel[s->cur] + s->poff, s->np_in, d, acc);
		FS3_LAUNCH_CHECK();
	}
	const float* g[4] = {s->U, s->V, s->W, s->phi};
	double sums[4] = {0, 0, 0, 0};
	for (int k = 0; k < 4; k++) {
		size_t q0, q1;
		if (k < 3) fs3i_p2g_range(s, k, &q0, &q1);
		else { q0 = (size_t)s->nx * s->ny * s->z0; q1 = (size_t)s->nx * s->ny * s->z1; }
		const size_t nb = (q1 - q0 + 255) / 256;
		k_check_grid<<<(unsigned)nb, 256, 0, s->stream>>>(g[k], q0, q1, acc + 3 + k, part + 2); FS3_LAUNCH_CHECK();
		k_partials_final<<<1, 256, 0, s->stream>>>(part + 2, nb, part); FS3_LAUNCH_CHECK();
		double h[2];
		FS3_CUDA(cudaMemcpyAsync(h, part, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
		FS3_CUDA(cudaStreamSynchronize(s->stream));
		sums[k] = h[1];
	}
	unsigned long long h[7];
	FS3_CUDA(cudaMemcpyAsync(h, acc, sizeof(h), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	out->particle_hash = h[0]; out->cell_hash = h[1]; out->particles = (long long)h[2];
	for (int k = 0; k < 4; k++) { out->grid_hash[k] = h[3 + k]; out->grid_sum[k] = sums[k]; }
	return FS3_OK;
}

int fs3_stage_advect(fs3_sim* s, float dt) { ENTER(s); SINGLE(s); return fs3i_advect(s, dt); }
int fs3_stage_bin(fs3_sim* s) { ENTER(s); SINGLE(s); return fs3i_bin(s); }
int fs3_stage_levelset(fs3_sim* s) { ENTER(s); SINGLE(s); return fs3i_levelset(s); }

// The stand-alone stage leaves its result in the live grids (as the reference does), so copy old -> cur.
int fs3_stage_p2g(fs3_sim* s) {
	ENTER(s);
	SINGLE(s);
	int rc = fs3i_p2g(s);
	if (rc) return rc;
	FS3_CUDA(cudaMemcpyAsync(s->U, s->oU, s->nU * 4, cudaMemcpyDeviceToDevice, s->stream));
	FS3_CUDA(cudaMemcpyAsync(s->V, s->oV, s->nV * 4, cudaMemcpyDeviceToDevice, s->stream));
	FS3_CUDA(cudaMemcpyAsync(s->W, s->oW, s->nW * 4, cudaMemcpyDeviceToDevice, s->stream));
	return FS3_OK;
}
int fs3_stage_bodyforces(fs3_sim* s, float dt) { ENTER(s); SINGLE(s); return fs3i_bodyforces(s, dt); }
int fs3_stage_project(fs3_sim* s, float dt) { ENTER(s); SINGLE(s); return fs3i_project(s, dt, false, false); }

int fs3_stage_g2p(fs3_sim* s, float dt, const float* oldU, const float* oldV, const float* oldW) {
	ENTER(s);
	SINGLE(s);
	if (!oldU || !oldV || !oldW) return FS3_ERR_INVALID;
	FS3_CUDA(cudaMemcpyAsync(s->oU, oldU, s->nU * 4, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaMemcpyAsync(s->oV, oldV, s->nV * 4, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaMemcpyAsync(s->oW, oldW, s->nW * 4, cudaMemcpyHostToDevice, s->stream));
	float alpha = flip_alpha(s, dt);
	int rc = fs3i_flip_combine(s, alpha);
	if (rc) return rc;
	rc = fs3i_g2p(s, alpha);
	if (rc) return rc;
	FS3_CUDA(cudaStreamSynchronize(s->stream));  // the host arrays may be released by the caller after return
	return FS3_OK;
}

int fs3_extrapolate(fs3_sim* s, float* grid, const unsigned char* valid, int xs, int ys, int zs) {
	ENTER(s);
	SINGLE(s);
	if (!grid || !valid || xs < 1 || ys < 1 || zs < 1) return FS3_ERR_INVALID;
	size_t n = (size_t)xs * ys * zs;
	size_t nmax = s->nW > s->nU ? s->nW : s->nU;
	if (s->nV > nmax) nmax = s->nV;
	if (n > nmax || xs + ys + zs + 1 > s->bucket_cap) return fail(s, FS3_ERR_INVALID, "grid larger than the handle's face grids");
	int rc = fs3i_staging(s, n * 5);
	if (rc) return rc;
	float* dg = (float*)s->stage_ptr;
	unsigned char* dv = (unsigned char*)s->stage_ptr + n * 4;
	FS3_CUDA(cudaMemcpyAsync(dg, grid, n * 4, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaMemcpyAsync(dv, valid, n, cudaMemcpyHostToDevice, s->stream));
	if ((rc = fs3i_cd_from_valid(s, dv, n))) return rc;
	if ((rc = fs3i_extrapolate(s, dg, xs, ys, zs))) return rc;
	FS3_CUDA(cudaMemcpyAsync(grid, dg, n * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}

int fs3_get_cell_indices(fs3_sim* s, int* cell) {
	ENTER(s);
	if (!cell) return FS3_ERR_INVALID;
	const long long total = fs3_num_particles(s);
	if (total <= 0) return FS3_OK;
	int rc = fs3i_staging(s, (size_t)total * 4);
	if (rc) return rc;
	if (s->world > 1) FS3_CUDA(cudaMemsetAsync(s->stage_ptr, 0xff, (size_t)total * 4, s->stream));   // -1 for particles of other ranks
	if ((rc = fs3i_cell_indices(s, (int*)s->stage_ptr))) return rc;
	FS3_CUDA(cudaMemcpyAsync(cell, s->stage_ptr, (size_t)total * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_cell_counts(fs3_sim* s, int* counts) {
	ENTER(s);
	if (!counts) return FS3_ERR_INVALID;
	SINGLE(s);
	int rc = fs3i_bin(s);
	if (rc) return rc;
	if ((rc = fs3i_staging(s, s->nC * 4))) return rc;
	k_cell_counts<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->cell_start, GridDims{s->nx, s->ny, s->nz}, (int*)s->stage_ptr); FS3_LAUNCH_CHECK();
	FS3_CUDA(cudaMemcpyAsync(counts, s->stage_ptr, s->nC * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_closest(fs3_sim* s, int* closest) {
	ENTER(s);
	if (!closest) return FS3_ERR_INVALID;
	if (!s->closest_valid) return fail(s, FS3_ERR_STATE, "no level set has been computed since the particles last moved");
	{ int rc = fs3i_field_wait(s); if (rc) return rc; }
	FS3_CUDA(cudaMemcpyAsync(closest, s->closest, s->nC * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
static int export_sys(fs3_sim* s, double* p, double* b, double* dg) {
	if (!s->system_valid) return fail(s, FS3_ERR_STATE, "no projection has been run yet");
	int rc = fs3i_staging(s, s->nC * 8 * 3);
	if (rc) return rc;
	double* d = (double*)s->stage_ptr;
	if ((rc = fs3i_export_pressure(s, p ? d : nullptr, b ? d + s->nC : nullptr, dg ? d + 2 * s->nC : nullptr))) return rc;
	if (p) FS3_CUDA(cudaMemcpyAsync(p, d, s->nC * 8, cudaMemcpyDeviceToHost, s->stream));
	if (b) FS3_CUDA(cudaMemcpyAsync(b, d + s->nC, s->nC * 8, cudaMemcpyDeviceToHost, s->stream));
	if (dg) FS3_CUDA(cudaMemcpyAsync(dg, d + 2 * s->nC, s->nC * 8, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_pressure(fs3_sim* s, double* p) { ENTER(s); if (!p) return FS3_ERR_INVALID; return export_sys(s, p, nullptr, nullptr); }
int fs3_get_rhs_diag(fs3_sim* s, double* b, double* diag) { ENTER(s); return export_sys(s, nullptr, b, diag); }

int fs3_get_sor_residuals(fs3_sim* s, double* l2, double* linf, int cap, int* count) {
	ENTER(s);
	if (count) *count = s->res_count;
	int n = s->res_count < cap ? s->res_count : cap;
	if (n <= 0) return FS3_OK;
	std::vector<double> h(2 * (size_t)n);
	FS3_CUDA(cudaMemcpyAsync(h.data(), s->res, sizeof(double) * 2 * n, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	for (int i = 0; i < n; i++) {
		if (l2) l2[i] = std::sqrt(h[2 * i]);
		if (linf) linf[i] = h[2 * i + 1];  // stored as the bit pattern of a non-negative double
	}
	return FS3_OK;
}

int fs3_sor_bench(fs3_sim* s, int iters, float* ms_out) {
	ENTER(s);
	if (iters < 1 || !ms_out) return FS3_ERR_INVALID;
	SINGLE(s);
	if (!s->system_valid) return fail(s, FS3_ERR_STATE, "run a projection first: the sweep needs b, diag and the masks");
	int rc = fs3i_clear_pressure(s);
	if (rc) return rc;
	int keep = s->prm.record_residuals;
	s->prm.record_residuals = 0;
	cudaEventRecord(s->ev[FS3_T_COUNT], s->stream);
	rc = fs3i_sor_loop(s, iters);
	cudaEventRecord(s->ev[FS3_T_COUNT + 1], s->stream);
	s->prm.record_residuals = keep;
	if (rc) return rc;
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	FS3_CUDA(cudaEventElapsedTime(ms_out, s->ev[FS3_T_COUNT], s->ev[FS3_T_COUNT + 1]));
	return FS3_OK;
}

int fs3_set_phi_all_fluid(fs3_sim* s) {
	ENTER(s);
	SINGLE(s);
	k_fill_f32<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->phi, s->nC, -1.0f); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

}  // extern "C"
