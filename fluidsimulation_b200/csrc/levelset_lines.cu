// "Line-sweep" level set (fs3_params.levelset_mode = 1): the reference's GPU formulation of
// ComputeLevelSet in place of the CPU path's 8 Gauss-Seidel sweeps -- phi cleared to +INF (Simulation.cpp:718-723), closest
// particle among the 27 neighbour cells (FX/gpComputeClosestParticleNeighbors.hlsl:40-112), then 24 one-directional line
// sweeps (FX/gpClosestParticlesSweep{Xm,Xp,Ym,Yp,Zm,Zp}.hlsl in the order of Simulation.cpp:744-753).  No sweep has a
// dependency that crosses lines, so there is no grid-long chain: a sweep is n^2 independent marches of n cells.  It is a
// DIFFERENT level set from the default mode's (more cells count as fluid, phi is closer to the exact distance), so it is
// opt-in and never used for parity with the reference's CPU class.
//
// Readings of what the shaders leave unspecified (identical to the CPU restatement the tests compare against):
// the solver's own cell rule roundf(n*X); particles of a cell in (octant, creation id) order, which only decides exact
// distance ties (first candidate scanned wins); every pass starts from the first frame's state -- every cell points at
// binned particle 0, the first particle of the lowest-numbered non-empty cell (the reference never clears its
// closest-particle texture); distance = sqrtf(dx*dx + dy*dy + dz*dz), unfused.  Kept quirks of the shader source: `+`
// sweeps take their first candidate from the cell at coordinate 0 of the line and visit coordinate 0; `-` sweeps start at 1.
//
// As in the default mode each cell carries its candidate as a record (n*X, n*Y, n*Z, creation id) in ls_rec, so a march
// never gathers a particle by index.
#include "fs3_internal.cuh"

__device__ __forceinline__ float gls_dist(float cx, float cy, float cz, float4 q, float radius) {
	float dx = __fsub_rn(cx, q.x), dy = __fsub_rn(cy, q.y), dz = __fsub_rn(cz, q.z);
	float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
	return __fsub_rn(sqrtf(s), radius);
}
__device__ __forceinline__ float4 gls_record(float4 p, GridDims d) {
	return make_float4(__fmul_rn(p.x, (float)d.nx), __fmul_rn(p.y, (float)d.ny), __fmul_rn(p.z, (float)d.nz), p.w);
}

// lowest-numbered non-empty cell -> *first (initialised to 0xffffffff)
__global__ void __launch_bounds__(256) k_gls_first_cell(const uint32_t* __restrict__ start, GridDims d, uint32_t* __restrict__ first) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	bool has = false;
	if (c < (size_t)d.nx * d.ny * d.nz) {
		int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
		FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) has = has || (e > b);
	}
	// one atomic per warp: the lowest lane that has particles holds the lowest cell of the warp
	const unsigned m = __ballot_sync(0xffffffffu, has);
	if (m && (threadIdx.x & 31) == __ffs(m) - 1) atomicMin(first, (uint32_t)c);
}
// binned particle 0 = first particle (octant, id order) of that cell
__global__ void k_gls_default_record(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, const uint32_t* __restrict__ first,
                                     float4* __restrict__ defrec) {
	const uint32_t c = *first;
	if (c == 0xffffffffu) { *defrec = make_float4(0.f, 0.f, 0.f, __int_as_float(-1)); return; }
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	bool done = false;
	FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) {
		if (!done && e > b) { *defrec = gls_record(pos[b], d); done = true; }
	}
}

// K7: thread per cell, the 27 neighbour cells z-outermost / x-innermost, each cell's four half-row runs in order.
__global__ void __launch_bounds__(256) k_gls_seed27(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, float radius,
                                                    const float4* __restrict__ defrec, float4* __restrict__ rec, float* __restrict__ phi) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= (size_t)d.nx * d.ny * d.nz) return;
	const int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	float4 best = *defrec;
	float bphi = __int_as_float(0x7f800000);   // +INF
	for (int dz = -1; dz <= 1; dz++) {
		const int Z = z + dz;
		if (Z < 0 || Z >= d.nz) continue;
		for (int dy = -1; dy <= 1; dy++) {
			const int Y = y + dy;
			if (Y < 0 || Y >= d.ny) continue;
			for (int dx = -1; dx <= 1; dx++) {
				const int X = x + dx;
				if (X < 0 || X >= d.nx) continue;
				FS3_FOR_CELL_RANGES(start, d, X, Y, Z, b, e) {
					for (uint32_t t = b; t < e; t++) {
						const float4 q = gls_record(__ldg(pos + t), d);
						const float k = gls_dist((float)x, (float)y, (float)z, q, radius);
						if (k < bphi) { bphi = k; best = q; }
					}
				}
			}
		}
	}
	rec[c] = best;
	phi[c] = bphi;
}

// K8-K13: one thread per line.  AXIS = direction of the march, PLUS = towards lower coordinates ("looking at +").
// The cell's phi and record do not depend on the carried candidate, so the next cell's are fetched one step ahead.
template <int AXIS, bool PLUS>
__global__ void __launch_bounds__(128) k_gls_sweep(float4* __restrict__ rec, float* __restrict__ phi, GridDims d, float radius) {
	const int n1 = AXIS == 0 ? d.ny : d.nx;                    // fastest-varying coordinate across threads
	const int n2 = AXIS == 2 ? d.ny : d.nz;
	const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= (size_t)n1 * n2) return;
	const int u = (int)(t % n1), v = (int)(t / n1);
	int p[3];
	if (AXIS == 0) { p[0] = 0; p[1] = u; p[2] = v; }
	else if (AXIS == 1) { p[0] = u; p[1] = 0; p[2] = v; }
	else { p[0] = u; p[1] = v; p[2] = 0; }
	const size_t stride = AXIS == 0 ? 1 : (AXIS == 1 ? (size_t)d.nx : (size_t)d.nx * d.ny);
	const size_t line = (size_t)p[0] + (size_t)d.nx * ((size_t)p[1] + (size_t)d.ny * p[2]);
	const int len = AXIS == 0 ? d.nx : (AXIS == 1 ? d.ny : d.nz);
	float4 last = rec[line];                                   // the cell at coordinate 0, in both directions
	int k = PLUS ? len - 1 : 1;
	const int step = PLUS ? -1 : 1;
	if (len < 2 && !PLUS) return;
	float4 r_next = rec[line + (size_t)k * stride];
	float f_next = phi[line + (size_t)k * stride];
	for (; PLUS ? k >= 0 : k < len; k += step) {
		const size_t c = line + (size_t)k * stride;
		const float4 r = r_next;
		const float f = f_next;
		const int kn = k + step;
		if (PLUS ? kn >= 0 : kn < len) { r_next = rec[line + (size_t)kn * stride]; f_next = phi[line + (size_t)kn * stride]; }
		const float cx = AXIS == 0 ? (float)k : (float)p[0], cy = AXIS == 1 ? (float)k : (float)p[1], cz = AXIS == 2 ? (float)k : (float)p[2];
		const float dist = gls_dist(cx, cy, cz, last, radius);
		if (dist < f) { phi[c] = dist; rec[c] = last; }
		else last = r;
	}
}

__global__ void __launch_bounds__(256) k_gls_closest(const float4* __restrict__ rec, int* __restrict__ closest, size_t n) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c < n) closest[c] = __float_as_int(rec[c].w);
}

template <int AXIS, bool PLUS> static void launch_sweep(fs3_sim* s, cudaStream_t st) {
	const GridDims d{s->nx, s->ny, s->nz};
	const long long lines = AXIS == 0 ? (long long)d.ny * d.nz : (AXIS == 1 ? (long long)d.nx * d.nz : (long long)d.nx * d.ny);
	k_gls_sweep<AXIS, PLUS><<<cdiv(lines, 128), 128, 0, st>>>(s->ls_rec, s->phi, d, s->prm.particle_radius);
}

// ComputeLevelSet in the line-sweep formulation, on the handle's stream; needs the sorted particles (runs the bin first).
int fs3i_levelset_lines(fs3_sim* s) {
	if (s->world > 1) { s->err = "the line-sweep level set is not available on a sharded handle"; return FS3_ERR_STATE; }
	if (s->np <= 0) { s->err = "the line-sweep level set needs at least one particle"; return FS3_ERR_STATE; }
	int rc = fs3i_bin(s);
	if (rc) return rc;
	cudaStream_t st = s->stream;
	const GridDims d{s->nx, s->ny, s->nz};
	if ((rc = fs3i_staging(s, 32))) return rc;
	uint32_t* first = (uint32_t*)s->stage_ptr;
	float4* defrec = (float4*)((char*)s->stage_ptr + 16);
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SEED], st);
	FS3_CUDA(cudaMemsetAsync(first, 0xff, 4, st));
	k_gls_first_cell<<<cdiv((long long)s->nC, 256), 256, 0, st>>>(s->cell_start, d, first); FS3_LAUNCH_CHECK();
	k_gls_default_record<<<1, 1, 0, st>>>(s->pos[s->cur], s->cell_start, d, first, defrec); FS3_LAUNCH_CHECK();
	k_gls_seed27<<<cdiv((long long)s->nC, 256), 256, 0, st>>>(s->pos[s->cur], s->cell_start, d, s->prm.particle_radius, defrec, s->ls_rec, s->phi); FS3_LAUNCH_CHECK();
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SWEEP], st);
	// 0 1 2 3 4 5 = x- x+ y- y+ z- z+ (Simulation.cpp:744-753)
	static const int dirs[24] = {0, 2, 4, 1, 2, 4, 0, 3, 4, 1, 3, 4, 0, 2, 5, 1, 2, 5, 0, 3, 5, 1, 3, 5};
	for (int k = 0; k < 24; k++) {
		switch (dirs[k]) {
			case 0: launch_sweep<0, false>(s, st); break;
			case 1: launch_sweep<0, true>(s, st); break;
			case 2: launch_sweep<1, false>(s, st); break;
			case 3: launch_sweep<1, true>(s, st); break;
			case 4: launch_sweep<2, false>(s, st); break;
			default: launch_sweep<2, true>(s, st); break;
		}
		FS3_LAUNCH_CHECK();
	}
	k_gls_closest<<<cdiv((long long)s->nC, 256), 256, 0, st>>>(s->ls_rec, s->closest, s->nC); FS3_LAUNCH_CHECK();
	s->closest_valid = true;
	s->ls_overlapped = false;
	return FS3_OK;
}
