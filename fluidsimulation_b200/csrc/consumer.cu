// Phi / binned-particle consumer interface: what the reference's GPU class hands to its
// renderer -- the public members m_gpPhiSRV, m_gpCountsSRV and m_gpBinnedParticles (Simulation.h:118-140), the phi blur
// it runs at the end of every step (K26, FX/gpBlur.hlsl:6-17, dispatched at Simulation.cpp:561-565) and the renderer's
// distance lookup `map()` (FX/Render.fx:77-115) -- as device buffers and kernels a downstream CUDA ray-marcher can use
// without a graphics API.  Nothing here is on Simulate()'s path: the solver's own phi is never modified.
//
// Where the reference is racy or hardware-defined, the deterministic reading is taken and said so:
//   * gpBlur updates the texture in place while neighbouring threads read it; here the 7-point mean reads the
//     un-blurred phi and writes a second buffer (Jacobi), out-of-range taps read 0 as D3D11 textures do;
//   * gpBinParticles orders the particles of a cell by InterlockedAdd arrival; here by octant, then creation id;
//   * Render.fx samples phi with the hardware's fixed-point trilinear filter; here the filter is explicit fp32
//     (x, then y, then z; a + (b - a) * w), clamp addressing, texel centres at (k + 0.5) / n.
#include "fs3_internal.cuh"

#define ENTER(s)                                                       \
	if (!(s)) return FS3_ERR_INVALID;                                  \
	do {                                                               \
		cudaError_t e0_ = cudaSetDevice((s)->device);                  \
		if (e0_ != cudaSuccess) { (s)->err = cudaGetErrorString(e0_); return FS3_ERR_CUDA; } \
	} while (0)

static int fail(fs3_sim* s, int code, const char* msg) { s->err = msg; return code; }

// ------------------------------------------------------------------------------------------------
// K26: out = (c + x+ + x- + y+ + y- + z+ + z-) / 7, summed left to right (gpBlur.hlsl:9-16).
// HBM-bound: 4 B read + 4 B written per cell; the six neighbour taps hit L1/L2.  A thread owns four
// consecutive x cells and marches BLUR_ZC planes up: per plane three 128-bit loads (next plane's centre row,
// the two y neighbours), two scalar x taps, one streaming 128-bit store; 3-D launch, no index divisions
// (the first version, one thread per 4 cells with a linear index, was issue-bound: 226 instructions per thread).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float blur7(float c, float xp, float xm, float yp, float ym, float zp, float zm) {
	float a = __fadd_rn(c, xp);
	a = __fadd_rn(a, xm); a = __fadd_rn(a, yp); a = __fadd_rn(a, ym); a = __fadd_rn(a, zp); a = __fadd_rn(a, zm);
	return __fdiv_rn(a, 7.0f);
}

#define BLUR_ZC 8   // planes a thread marches through: the z taps of plane k are the centre rows of planes k-1 / k+1, kept in registers
__global__ void __launch_bounds__(256) k_blur_phi_v4(const float* __restrict__ phi, float* __restrict__ out, int nx, int ny, int nz) {
	const int x4 = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y, z0 = blockIdx.z * BLUR_ZC;
	if (x4 >= (nx >> 2) || y >= ny) return;
	const size_t plane = (size_t)nx * ny;
	size_t base = 4 * (size_t)x4 + (size_t)nx * ((size_t)y + (size_t)ny * z0);
	const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
	const bool has_l = x4 > 0, has_r = 4 * x4 + 4 < nx, has_yp = y + 1 < ny, has_ym = y > 0;
	float4 zm = z0 > 0 ? __ldg(reinterpret_cast<const float4*>(phi + base - plane)) : zero;
	float4 c = __ldg(reinterpret_cast<const float4*>(phi + base));
	const int z1 = min(z0 + BLUR_ZC, nz);
	for (int z = z0; z < z1; z++, base += plane) {
		const float4 zp = z + 1 < nz ? __ldg(reinterpret_cast<const float4*>(phi + base + plane)) : zero;
		const float4 yp = has_yp ? __ldg(reinterpret_cast<const float4*>(phi + base + nx)) : zero;
		const float4 ym = has_ym ? __ldg(reinterpret_cast<const float4*>(phi + base - nx)) : zero;
		const float xl = has_l ? __ldg(phi + base - 1) : 0.f;
		const float xr = has_r ? __ldg(phi + base + 4) : 0.f;
		float4 r;
		r.x = blur7(c.x, c.y, xl, yp.x, ym.x, zp.x, zm.x);
		r.y = blur7(c.y, c.z, c.x, yp.y, ym.y, zp.y, zm.y);
		r.z = blur7(c.z, c.w, c.y, yp.z, ym.z, zp.z, zm.z);
		r.w = blur7(c.w, xr, c.z, yp.w, ym.w, zp.w, zm.w);
		__stcs(reinterpret_cast<float4*>(out + base), r);   // written once, read by another kernel: do not displace phi in L2
		zm = c; c = zp;
	}
}

__global__ void __launch_bounds__(256) k_blur_phi_scalar(const float* __restrict__ phi, float* __restrict__ out, int nx, int ny, int nz) {
	const size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= (size_t)nx * ny * nz) return;
	const int x = (int)(q % nx), y = (int)((q / nx) % ny), z = (int)(q / ((size_t)nx * ny));
	const size_t plane = (size_t)nx * ny;
	out[q] = blur7(__ldg(phi + q), x + 1 < nx ? __ldg(phi + q + 1) : 0.f, x > 0 ? __ldg(phi + q - 1) : 0.f, y + 1 < ny ? __ldg(phi + q + nx) : 0.f,
	               y > 0 ? __ldg(phi + q - nx) : 0.f, z + 1 < nz ? __ldg(phi + q + plane) : 0.f, z > 0 ? __ldg(phi + q - plane) : 0.f);
}

// ------------------------------------------------------------------------------------------------
// Render.fx:77-115 `map(p)`: distance to the fluid surface, in metres, at a point of the unit box.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float quintic(float f) {   // f*f*f*(f*(f*6 - 15) + 10), Render.fx:85, never fused
	float a = __fmul_rn(__fmul_rn(f, f), f);
	float b = __fadd_rn(__fmul_rn(f, __fsub_rn(__fmul_rn(f, 6.0f), 15.0f)), 10.0f);
	return __fmul_rn(a, b);
}
// one axis: unit-box coordinate -> (lower texel, upper texel, weight) of the clamped linear filter
__device__ __forceinline__ void map_axis(float p, float n, float inv, int last, bool smooth, int& i0, int& i1, float& w) {
	float mp = __fadd_rn(__fmul_rn(n, p), 0.5f);                   // Render.fx:81
	float i = floorf(mp);
	float f = __fsub_rn(mp, i);
	if (smooth) f = quintic(f);                                    // x and z only (Render.fx:85)
	mp = __fadd_rn(i, f);                                          // Render.fx:111
	float u = __fmul_rn(inv, __fsub_rn(mp, 0.5f));                 // Render.fx:113, normalised texture coordinate
	float t = __fsub_rn(__fmul_rn(u, n), 0.5f);                    // texel space: centre of texel k at (k + 0.5) / n
	float fl = floorf(t);
	w = __fsub_rn(t, fl);
	// clamp in float first: a far-away sample point must not overflow the int conversion
	fl = fl < -1.0f ? -1.0f : (fl > (float)last ? (float)last : fl);
	int k = (int)fl;
	i0 = k < 0 ? 0 : k;
	i1 = k + 1 > last ? last : k + 1;
}
__global__ void __launch_bounds__(256) k_sample_phi(const float* __restrict__ phi, GridDims d, const float* __restrict__ xyz, long long count, float* __restrict__ out) {
	long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count) return;
	const float inv = __fdiv_rn(1.0f, (float)d.nx), invy = __fdiv_rn(1.0f, (float)d.ny), invz = __fdiv_rn(1.0f, (float)d.nz);   // Render.fx:533
	int x0, x1, y0, y1, z0, z1;
	float wx, wy, wz;
	map_axis(xyz[3 * i + 0], (float)d.nx, inv, d.nx - 1, true, x0, x1, wx);
	map_axis(xyz[3 * i + 1], (float)d.ny, invy, d.ny - 1, false, y0, y1, wy);
	map_axis(xyz[3 * i + 2], (float)d.nz, invz, d.nz - 1, true, z0, z1, wz);
	const size_t nx = d.nx, pl = (size_t)d.nx * d.ny;
	const float* r00 = phi + nx * y0 + pl * z0;
	const float* r10 = phi + nx * y1 + pl * z0;
	const float* r01 = phi + nx * y0 + pl * z1;
	const float* r11 = phi + nx * y1 + pl * z1;
	float a00 = fs3_lerp(__ldg(r00 + x0), __ldg(r00 + x1), wx);
	float a10 = fs3_lerp(__ldg(r10 + x0), __ldg(r10 + x1), wx);
	float a01 = fs3_lerp(__ldg(r01 + x0), __ldg(r01 + x1), wx);
	float a11 = fs3_lerp(__ldg(r11 + x0), __ldg(r11 + x1), wx);
	float v = fs3_lerp(fs3_lerp(a00, a10, wy), fs3_lerp(a01, a11, wy), wz);
	out[i] = __fmul_rn(v, inv);                                    // Render.fx:115: phi is in cell units, the renderer marches in metres
}

// ------------------------------------------------------------------------------------------------
// Cell-major export of the sorted particles: offsets[c] .. offsets[c+1] are the records of cell c
// (gpBinParticles.hlsl:13-31 writes the same thing; the prefix sum there is the host loop
// Simulation.cpp:660-686).  The live particles are sorted by HALF-cell, so a cell is four runs.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_export_counts(const uint32_t* __restrict__ start, GridDims d, uint32_t* __restrict__ out) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	const size_t nc = (size_t)d.nx * d.ny * d.nz;
	if (q > nc) return;
	uint32_t n = 0;
	if (q < nc) {
		int x = (int)(q % d.nx), y = (int)((q / d.nx) % d.ny), z = (int)(q / ((size_t)d.nx * d.ny));
		FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) n += e - b;
	}
	out[q] = n;   // [nc] = 0, so that the exclusive scan leaves the total there
}
__global__ void __launch_bounds__(256) k_export_binned(const uint32_t* __restrict__ start, GridDims d, const float4* __restrict__ pos, const float4* __restrict__ vel,
                                                       const uint32_t* __restrict__ offsets, float2* __restrict__ out) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q >= (size_t)d.nx * d.ny * d.nz) return;
	int x = (int)(q % d.nx), y = (int)((q / d.nx) % d.ny), z = (int)(q / ((size_t)d.nx * d.ny));
	size_t o = offsets[q];
	FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) {
		for (uint32_t j = b; j < e; j++, o++) {
			float4 p = __ldg(pos + j), v = __ldg(vel + j);
			out[3 * o + 0] = make_float2(p.x, p.y);
			out[3 * o + 1] = make_float2(p.z, v.x);
			out[3 * o + 2] = make_float2(v.y, v.z);
		}
	}
}

static int ensure(fs3_sim* s, void** p, size_t* have, size_t bytes) {
	if (*p && *have >= bytes) return FS3_OK;
	if (*p) cudaFree(*p);
	*p = nullptr; *have = 0;
	FS3_CUDA(cudaMalloc(p, bytes));
	*have = bytes;
	return FS3_OK;
}

static int blur(fs3_sim* s) {
	int rc = ensure(s, (void**)&s->phi_blur, &s->phi_blur_bytes, s->nC * 4);
	if (rc) return rc;
	if ((s->nx & 3) == 0) {
		dim3 grid(cdiv(s->nx >> 2, 32), cdiv(s->ny, 8), cdiv(s->nz, BLUR_ZC));
		k_blur_phi_v4<<<grid, dim3(32, 8, 1), 0, s->stream>>>(s->phi, s->phi_blur, s->nx, s->ny, s->nz); FS3_LAUNCH_CHECK();
	}
	else { k_blur_phi_scalar<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->phi, s->phi_blur, s->nx, s->ny, s->nz); FS3_LAUNCH_CHECK(); }
	s->phi_blur_valid = true;
	return FS3_OK;
}

static int export_binned(fs3_sim* s) {
	if (s->world > 1) return fail(s, FS3_ERR_STATE, "the binned-particle export is not available on a sharded handle");
	int rc = fs3i_bin(s);
	if (rc) return rc;
	if ((rc = ensure(s, (void**)&s->bin_offsets, &s->bin_offsets_bytes, (s->nC + 1) * 4))) return rc;
	if ((rc = ensure(s, (void**)&s->binned, &s->binned_bytes, (size_t)(s->np > 0 ? s->np : 1) * sizeof(fs3_particle)))) return rc;
	if ((rc = fs3i_staging(s, (s->nC + 1) * 4))) return rc;
	GridDims d{s->nx, s->ny, s->nz};
	k_export_counts<<<cdiv((long long)s->nC + 1, 256), 256, 0, s->stream>>>(s->cell_start, d, (uint32_t*)s->stage_ptr); FS3_LAUNCH_CHECK();
	if ((rc = fs3i_scan_u32(s, (const uint32_t*)s->stage_ptr, s->bin_offsets, s->nC + 1))) return rc;
	if (s->np > 0) {
		k_export_binned<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->cell_start, d, s->pos[s->cur], s->vel[s->cur], s->bin_offsets, (float2*)s->binned);
		FS3_LAUNCH_CHECK();
	}
	return FS3_OK;
}

void fs3i_consumer_free(fs3_sim* s) {
	if (s->phi_blur) cudaFree(s->phi_blur);
	if (s->bin_offsets) cudaFree(s->bin_offsets);
	if (s->binned) cudaFree(s->binned);
	s->phi_blur = nullptr; s->bin_offsets = nullptr; s->binned = nullptr;
	s->phi_blur_bytes = s->bin_offsets_bytes = s->binned_bytes = 0;
}

extern "C" {

int fs3_phi_device(fs3_sim* s, const float** dev_phi) {
	if (!s || !dev_phi) return FS3_ERR_INVALID;
	*dev_phi = s->phi;
	return FS3_OK;
}

int fs3_blur_phi(fs3_sim* s) { ENTER(s); return blur(s); }

int fs3_phi_blurred_device(fs3_sim* s, const float** dev_phi) {
	if (!s || !dev_phi) return FS3_ERR_INVALID;
	if (!s->phi_blur_valid) return fail(s, FS3_ERR_STATE, "fs3_blur_phi has not been called");
	*dev_phi = s->phi_blur;
	return FS3_OK;
}

int fs3_get_phi_blurred(fs3_sim* s, float* host) {
	ENTER(s);
	if (!host) return FS3_ERR_INVALID;
	if (!s->phi_blur_valid) return fail(s, FS3_ERR_STATE, "fs3_blur_phi has not been called");
	FS3_CUDA(cudaMemcpyAsync(host, s->phi_blur, s->nC * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}

int fs3_export_binned(fs3_sim* s, const fs3_particle** dev_binned, const unsigned** dev_offsets) {
	ENTER(s);
	int rc = export_binned(s);
	if (rc) return rc;
	if (dev_binned) *dev_binned = (const fs3_particle*)s->binned;
	if (dev_offsets) *dev_offsets = s->bin_offsets;
	return FS3_OK;
}

int fs3_get_binned(fs3_sim* s, fs3_particle* host_binned, unsigned* host_offsets) {
	ENTER(s);
	int rc = export_binned(s);
	if (rc) return rc;
	if (host_binned && s->np > 0) FS3_CUDA(cudaMemcpyAsync(host_binned, s->binned, (size_t)s->np * sizeof(fs3_particle), cudaMemcpyDeviceToHost, s->stream));
	if (host_offsets) FS3_CUDA(cudaMemcpyAsync(host_offsets, s->bin_offsets, (s->nC + 1) * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}

int fs3_sample_phi(fs3_sim* s, const float* xyz, long long count, int blurred, float* out) {
	ENTER(s);
	if (count < 0 || (count > 0 && (!xyz || !out))) return FS3_ERR_INVALID;
	if (blurred && !s->phi_blur_valid) return fail(s, FS3_ERR_STATE, "fs3_blur_phi has not been called");
	if (count == 0) return FS3_OK;
	int rc = fs3i_staging(s, (size_t)count * 16);
	if (rc) return rc;
	float* dxyz = (float*)s->stage_ptr;
	float* dout = dxyz + 3 * (size_t)count;
	FS3_CUDA(cudaMemcpyAsync(dxyz, xyz, (size_t)count * 12, cudaMemcpyHostToDevice, s->stream));
	k_sample_phi<<<cdiv(count, 256), 256, 0, s->stream>>>(blurred ? s->phi_blur : s->phi, GridDims{s->nx, s->ny, s->nz}, dxyz, count, dout); FS3_LAUNCH_CHECK();
	FS3_CUDA(cudaMemcpyAsync(out, dout, (size_t)count * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}

int fs3_sample_phi_device(fs3_sim* s, const float* dev_xyz, long long count, int blurred, float* dev_out) {
	ENTER(s);
	if (count < 0 || (count > 0 && (!dev_xyz || !dev_out))) return FS3_ERR_INVALID;
	if (blurred && !s->phi_blur_valid) return fail(s, FS3_ERR_STATE, "fs3_blur_phi has not been called");
	if (count == 0) return FS3_OK;
	k_sample_phi<<<cdiv(count, 256), 256, 0, s->stream>>>(blurred ? s->phi_blur : s->phi, GridDims{s->nx, s->ny, s->nz}, dev_xyz, count, dev_out); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

}  // extern "C"
