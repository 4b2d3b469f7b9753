// ComputeLevelSet (S3D.cpp:255-420) on sm_100a: per-cell seed from the cell-sorted particles, then the
// reference's 8 in-place Gauss-Seidel sweeps reproduced EXACTLY as block wavefronts.
//
// In the sweep with direction signs (sx,sy,sz) a cell only reads its three upwind neighbours, after
// they were finalised in the same sweep, and its three offers stay in x,y,z order inside one thread.
// Any schedule that respects that partial order gives the reference's result bit for bit, so tiles of
// 8x8x8 cells are processed in tile-hyperplane order (one launch per tile plane) and inside a tile the
// 22 cell hyperplanes run between __syncthreads() with the tile + its upwind halo held in shared memory.
// Each cell carries its closest particle as (n*X, n*Y, n*Z, id): the candidate's position travels with
// the index, so the dependent chain contains no global gather, and phi is recomputed from it
// (ptDistance is symmetric in its arguments, S3D.cpp:234-236).
#include "fs3_internal.cuh"

#define LS_T 8  // tile edge

__device__ __forceinline__ float ls_dist(float cx, float cy, float cz, float4 q, float radius) {
	float dx = __fsub_rn(cx, q.x), dy = __fsub_rn(cy, q.y), dz = __fsub_rn(cz, q.z);
	float s = __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
	return __fsub_rn(sqrtf(s), radius);
}

// Seed (S3D.cpp:281-300): minimum kernel distance over the particles of the cell; the segment is in
// creation order so the reference's strict `phi > k` keeps the lowest index on ties.
__global__ void __launch_bounds__(256) k_ls_seed(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, float radius,
                                                 float4* __restrict__ rec) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	size_t nc = (size_t)d.nx * d.ny * d.nz;
	if (c >= nc) return;
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	uint32_t b = start[c], e = start[c + 1];
	float4 best = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	float bk = 0.f;
	bool have = false;
	for (uint32_t t = b; t < e; t++) {
		float4 p = __ldg(pos + t);
		float4 q = make_float4(__fmul_rn(p.x, (float)d.nx), __fmul_rn(p.y, (float)d.ny), __fmul_rn(p.z, (float)d.nz), p.w);
		float k = ls_dist((float)x, (float)y, (float)z, q, radius);
		if (!have || bk > k) { have = true; bk = k; best = q; }
	}
	rec[c] = best;
}

// One tile hyperplane of one sweep.  Sweep space: X' = x (dir +) or nx-1-x (dir -), same for Y', Z';
// tiles and wavefronts are defined in sweep space, only addresses use the flip.
__global__ void __launch_bounds__(LS_T* LS_T* LS_T) k_ls_sweep_plane(float4* __restrict__ rec, GridDims d, int flipx, int flipy, int flipz,
                                                                    int wplane, int ntx, float radius) {
	__shared__ float4 sm[(LS_T + 1) * (LS_T + 1) * (LS_T + 1)];
	int Ty = blockIdx.x, Tz = blockIdx.y;
	int Tx = wplane - Ty - Tz;
	if (Tx < 0 || Tx >= ntx) return;
	int tid = threadIdx.x;
	int lx = tid % LS_T, ly = (tid / LS_T) % LS_T, lz = tid / (LS_T * LS_T);
	const int P = LS_T + 1;
#define SM(i, j, k) sm[((i) + 1) + P * (((j) + 1) + P * ((k) + 1))]
	int X = Tx * LS_T + lx, Y = Ty * LS_T + ly, Z = Tz * LS_T + lz;  // sweep-space coordinates
	bool inside = X < d.nx && Y < d.ny && Z < d.nz;
	int gx = flipx ? d.nx - 1 - X : X, gy = flipy ? d.ny - 1 - Y : Y, gz = flipz ? d.nz - 1 - Z : Z;
	size_t gi = (size_t)gx + (size_t)d.nx * ((size_t)gy + (size_t)d.ny * gz);
	float4 mine = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	if (inside) mine = rec[gi];
	SM(lx, ly, lz) = mine;
	// upwind halo faces: 3 * 64 records, loaded by the first 192 threads
	if (tid < 3 * LS_T * LS_T) {
		int f = tid / (LS_T * LS_T), a = tid % LS_T, b = (tid / LS_T) % LS_T;
		int hx = f == 0 ? -1 : a, hy = f == 1 ? -1 : (f == 0 ? a : b), hz = f == 2 ? -1 : b;
		int HX = Tx * LS_T + hx, HY = Ty * LS_T + hy, HZ = Tz * LS_T + hz;
		float4 h = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
		if (HX >= 0 && HY >= 0 && HZ >= 0 && HX < d.nx && HY < d.ny && HZ < d.nz) {
			int qx = flipx ? d.nx - 1 - HX : HX, qy = flipy ? d.ny - 1 - HY : HY, qz = flipz ? d.nz - 1 - HZ : HZ;
			h = rec[(size_t)qx + (size_t)d.nx * ((size_t)qy + (size_t)d.ny * qz)];
		}
		SM(hx, hy, hz) = h;
	}
	__syncthreads();
	float cx = (float)gx, cy = (float)gy, cz = (float)gz;
	int myid = __float_as_int(mine.w);
	float myphi = myid >= 0 ? ls_dist(cx, cy, cz, mine, radius) : 0.f;
	bool changed = false;
	int mystep = lx + ly + lz;
	for (int step = 0; step <= 3 * (LS_T - 1); step++) {
		if (step == mystep && inside) {
			// clsInner on the upwind x-, y-, z-neighbour in that order (S3D.cpp:310-315 and siblings):
			// index <= 0 is never propagated (:242); strict `<` keeps the incumbent (:248).
#pragma unroll
			for (int a = 0; a < 3; a++) {
				bool has = a == 0 ? X != 0 : (a == 1 ? Y != 0 : Z != 0);
				if (!has) continue;
				float4 cand = a == 0 ? SM(lx - 1, ly, lz) : (a == 1 ? SM(lx, ly - 1, lz) : SM(lx, ly, lz - 1));
				if (__float_as_int(cand.w) > 0) {
					float dd = ls_dist(cx, cy, cz, cand, radius);
					if (myid < 0 || dd < myphi) { mine = cand; myid = __float_as_int(cand.w); myphi = dd; changed = true; }
				}
			}
			SM(lx, ly, lz) = mine;
		}
		__syncthreads();
	}
	if (changed) rec[gi] = mine;
#undef SM
}

// phi and closestParticles[] from the records; cells never reached keep their previous phi, as in
// the reference where m_Phi is only written together with closestParticles (S3D.cpp:248-251,296-299).
__global__ void __launch_bounds__(256) k_ls_finalize(const float4* __restrict__ rec, GridDims d, float radius, float* __restrict__ phi,
                                                     int* __restrict__ closest) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	size_t nc = (size_t)d.nx * d.ny * d.nz;
	if (c >= nc) return;
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	float4 r = rec[c];
	int id = __float_as_int(r.w);
	closest[c] = id;
	if (id >= 0) phi[c] = ls_dist((float)x, (float)y, (float)z, r, radius);
}

int fs3i_levelset(fs3_sim* s) {
	int rc = fs3i_bin(s);
	if (rc) return rc;
	GridDims d{s->nx, s->ny, s->nz};
	float radius = s->prm.particle_radius;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SEED], s->stream);
	k_ls_seed<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->pos[s->cur], s->cell_start, d, radius, s->ls_rec);
	FS3_LAUNCH_CHECK();
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_LEVELSET_SWEEP], s->stream);
	int ntx = cdiv(d.nx, LS_T), nty = cdiv(d.ny, LS_T), ntz = cdiv(d.nz, LS_T);
	dim3 grid(nty, ntz);
	for (int m = 0; m < 8; m++) {  // sweep order of S3D.cpp:306-416: bit0 -> x descending, bit1 -> y, bit2 -> z
		for (int w = 0; w <= ntx + nty + ntz - 3; w++) {
			k_ls_sweep_plane<<<grid, LS_T * LS_T * LS_T, 0, s->stream>>>(s->ls_rec, d, m & 1, (m >> 1) & 1, (m >> 2) & 1, w, ntx, radius);
			FS3_LAUNCH_CHECK();
		}
	}
	k_ls_finalize<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->ls_rec, d, radius, s->phi, s->closest);
	FS3_LAUNCH_CHECK();
	s->closest_valid = true;
	return FS3_OK;
}
