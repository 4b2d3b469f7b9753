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

extern "C" {

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
