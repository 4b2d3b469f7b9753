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

// Seed (S3D.cpp:281-300): minimum kernel distance over the particles of the cell (its 8 half-cells); the
// reference's strict `phi > k` in index order keeps the lowest index on ties.
__global__ void __launch_bounds__(256) k_ls_seed(const float4* __restrict__ pos, const uint32_t* __restrict__ start, GridDims d, float radius,
                                                 float4* __restrict__ rec) {
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	size_t nc = (size_t)d.nx * d.ny * d.nz;
	if (c >= nc) return;
	int x = (int)(c % d.nx), y = (int)((c / d.nx) % d.ny), z = (int)(c / ((size_t)d.nx * d.ny));
	float4 best = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
	float bk = 0.f;
	int bid = -1;
	FS3_FOR_CELL_RANGES(start, d, x, y, z, b, e) {
		for (uint32_t t = b; t < e; t++) {
			float4 p = __ldg(pos + t);
			float4 q = make_float4(__fmul_rn(p.x, (float)d.nx), __fmul_rn(p.y, (float)d.ny), __fmul_rn(p.z, (float)d.nz), p.w);
			float k = ls_dist((float)x, (float)y, (float)z, q, radius);
			int id = __float_as_int(p.w);
			// reference: particles in index order, replace on strictly smaller distance (S3D.cpp:296) == min by (k, id)
			if (bid < 0 || k < bk || (k == bk && id < bid)) { bk = k; bid = id; best = q; }
		}
	}
	rec[c] = best;
}

// All 8 sweeps in ONE persistent launch, scheduled as a dataflow over 8x8x8 tiles.
// Tiles are a fixed partition of physical space.  Work items (sweep m, tile) are handed out through a
// ticket counter in sweep-major, tile-hyperplane order; a block that holds an item waits (polling the
// per-tile completion epochs in L2) until
//   * the tile itself and its 6 face neighbours have completed sweep m-1 (RAW on its own cells, WAR on
//     the halo cells its neighbours read in sweep m-1), and
//   * its 3 upwind neighbours have completed sweep m (RAW on the halo),
// all of which hold lower tickets, so some running block can always make progress (no co-residency
// assumption, no grid-wide barrier).  Inside a tile the 22 cell hyperplanes run between
// __syncthreads() with the tile and a one-cell halo in shared memory.  Consecutive sweeps overlap
// wherever the dependencies allow.  Records are read/written with L2-only accesses (ld.cg / st.cg):
// L1 is not coherent across SMs.
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }

__global__ void __launch_bounds__(LS_T* LS_T* LS_T) k_ls_sweep_dataflow(float4* __restrict__ rec, GridDims d, float radius,
                                                                       const uint32_t* __restrict__ order, uint32_t* done,
                                                                       uint32_t* ticket, int ntiles, int ntx, int nty, int ntz,
                                                                       uint32_t epoch0, int* __restrict__ flags) {
	const int P = LS_T + 2;
	__shared__ float4 sm[P * P * P];
	__shared__ uint32_t s_ticket;
	__shared__ int s_abort;
#define SM(i, j, k) sm[((i) + 1) + P * (((j) + 1) + P * ((k) + 1))]
	const int tid = threadIdx.x;
	const int lx = tid % LS_T, ly = (tid / LS_T) % LS_T, lz = tid / (LS_T * LS_T);
	if (tid == 0) s_abort = 0;
	for (;;) {
		__syncthreads();  // previous item fully retired (smem reuse, s_ticket reuse)
		if (tid == 0) s_ticket = atomicAdd(ticket, 1u);
		__syncthreads();
		const uint32_t t = s_ticket;
		if (t >= 8u * (uint32_t)ntiles || s_abort) break;
		const int m = (int)(t / (uint32_t)ntiles);
		const uint32_t packed = order[t % (uint32_t)ntiles];
		const int fx = m & 1, fy = (m >> 1) & 1, fz = (m >> 2) & 1;   // sweep order of S3D.cpp:306-416
		// tile in sweep space -> physical tile
		const int Tx = fx ? ntx - 1 - (int)(packed & 1023u) : (int)(packed & 1023u);
		const int Ty = fy ? nty - 1 - (int)((packed >> 10) & 1023u) : (int)((packed >> 10) & 1023u);
		const int Tz = fz ? ntz - 1 - (int)(packed >> 20) : (int)(packed >> 20);
		const int dx = fx ? -1 : 1, dy = fy ? -1 : 1, dz = fz ? -1 : 1;   // sweep direction per axis
		const uint32_t E = epoch0 + (uint32_t)m + 1u;
		// ---- wait for the dependencies: threads 0..6 poll one tile each ------------------------------
		if (tid < 7) {
			int qx = Tx, qy = Ty, qz = Tz;
			uint32_t need = E - 1u;
			if (tid == 1) qx -= 1; else if (tid == 2) qx += 1; else if (tid == 3) qy -= 1; else if (tid == 4) qy += 1;
			else if (tid == 5) qz -= 1; else if (tid == 6) qz += 1;
			if ((tid == 1 && dx > 0) || (tid == 2 && dx < 0) || (tid == 3 && dy > 0) || (tid == 4 && dy < 0) || (tid == 5 && dz > 0) ||
			    (tid == 6 && dz < 0))
				need = E;   // upwind neighbour: must have finished THIS sweep
			if (qx >= 0 && qx < ntx && qy >= 0 && qy < nty && qz >= 0 && qz < ntz) {
				const uint32_t* f = done + ((size_t)qx + (size_t)ntx * ((size_t)qy + (size_t)nty * qz));
				long long t0 = clock64();
				while ((int32_t)(ld_volatile_u32(f) - need) < 0) {
					__nanosleep(40);
					if (clock64() - t0 > 4000000000LL) { flags[1] = 1; s_abort = 1; break; }   // ~2 s: never hang the GPU
				}
			}
			__threadfence();
		}
		__syncthreads();
		if (s_abort) break;
		// ---- load the tile and its three upwind halo faces ---------------------------------------------
		const int x0 = Tx * LS_T, y0 = Ty * LS_T, z0 = Tz * LS_T;
		const int wx = min(LS_T, d.nx - x0), wy = min(LS_T, d.ny - y0), wz = min(LS_T, d.nz - z0);
		const int gx = x0 + lx, gy = y0 + ly, gz = z0 + lz;
		const bool inside = lx < wx && ly < wy && lz < wz;
		const size_t gi = (size_t)gx + (size_t)d.nx * ((size_t)gy + (size_t)d.ny * gz);
		float4 mine = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
		if (inside) mine = __ldcg(rec + gi);
		SM(lx, ly, lz) = mine;
		if (tid < 3 * LS_T * LS_T) {
			const int f = tid / (LS_T * LS_T), a = tid % LS_T, b = (tid / LS_T) % LS_T;
			// face f: the plane just upwind of the tile on axis f
			int hx = f == 0 ? (dx > 0 ? -1 : wx) : a;
			int hy = f == 1 ? (dy > 0 ? -1 : wy) : (f == 0 ? a : b);
			int hz = f == 2 ? (dz > 0 ? -1 : wz) : b;
			const int qx = x0 + hx, qy = y0 + hy, qz = z0 + hz;
			float4 h = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
			if (qx >= 0 && qy >= 0 && qz >= 0 && qx < d.nx && qy < d.ny && qz < d.nz)
				h = __ldcg(rec + ((size_t)qx + (size_t)d.nx * ((size_t)qy + (size_t)d.ny * qz)));
			if (hx <= LS_T && hy <= LS_T && hz <= LS_T) SM(hx, hy, hz) = h;
		}
		__syncthreads();
		// ---- 22 cell hyperplanes inside the tile ----------------------------------------------------------
		const float cx = (float)gx, cy = (float)gy, cz = (float)gz;
		int myid = __float_as_int(mine.w);
		float myphi = myid >= 0 ? ls_dist(cx, cy, cz, mine, radius) : 0.f;
		bool changed = false;
		const int mystep = (dx > 0 ? lx : wx - 1 - lx) + (dy > 0 ? ly : wy - 1 - ly) + (dz > 0 ? lz : wz - 1 - lz);
		const bool hasx = dx > 0 ? gx != 0 : gx != d.nx - 1, hasy = dy > 0 ? gy != 0 : gy != d.ny - 1, hasz = dz > 0 ? gz != 0 : gz != d.nz - 1;
		for (int step = 0; step <= 3 * (LS_T - 1); step++) {
			if (step == mystep && inside) {
				// clsInner on the upwind x-, y-, z-neighbour in that order (S3D.cpp:310-315 and siblings):
				// index <= 0 is never propagated (:242); strict `<` keeps the incumbent (:248).
#pragma unroll
				for (int a = 0; a < 3; a++) {
					const bool has = a == 0 ? hasx : (a == 1 ? hasy : hasz);
					if (!has) continue;
					const float4 cand = a == 0 ? SM(lx - dx, ly, lz) : (a == 1 ? SM(lx, ly - dy, lz) : SM(lx, ly, lz - dz));
					if (__float_as_int(cand.w) > 0) {
						const float dd = ls_dist(cx, cy, cz, cand, radius);
						if (myid < 0 || dd < myphi) { mine = cand; myid = __float_as_int(cand.w); myphi = dd; changed = true; }
					}
				}
				SM(lx, ly, lz) = mine;
			}
			__syncthreads();
		}
		if (changed) __stcg(rec + gi, mine);
		// ---- publish: all stores of the block, then the epoch -------------------------------------------
		__threadfence();
		__syncthreads();
		if (tid == 0) {
			*reinterpret_cast<volatile uint32_t*>(done + ((size_t)Tx + (size_t)ntx * ((size_t)Ty + (size_t)nty * Tz))) = E;
		}
	}
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
	FS3_CUDA(cudaMemsetAsync(s->ls_ticket, 0, sizeof(uint32_t), s->stream));
	k_ls_sweep_dataflow<<<s->ls_blocks, LS_T * LS_T * LS_T, 0, s->stream>>>(s->ls_rec, d, radius, s->ls_order, s->ls_done, s->ls_ticket,
	                                                                         s->ls_ntiles, ntx, nty, ntz, s->ls_epoch, s->dev_flags);
	FS3_LAUNCH_CHECK();
	s->ls_epoch += 8;
	k_ls_finalize<<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(s->ls_rec, d, radius, s->phi, s->closest);
	FS3_LAUNCH_CHECK();
	s->closest_valid = true;
	return FS3_OK;
}

// tile order table, completion epochs and the persistent grid size (called once from fs3_create)
#include <vector>
#include <algorithm>
int fs3i_levelset_init(fs3_sim* s) {
	int ntx = cdiv(s->nx, LS_T), nty = cdiv(s->ny, LS_T), ntz = cdiv(s->nz, LS_T);
	if (ntx > 1023 || nty > 1023 || ntz > 1023) { s->err = "grid too large for the level-set tile table"; return FS3_ERR_INVALID; }
	int nt = ntx * nty * ntz;
	std::vector<uint32_t> ord;
	ord.reserve(nt);
	for (int w = 0; w <= ntx + nty + ntz - 3; w++)
		for (int tz = 0; tz < ntz; tz++)
			for (int ty = 0; ty < nty; ty++) {
				int tx = w - ty - tz;
				if (tx >= 0 && tx < ntx) ord.push_back((uint32_t)tx | ((uint32_t)ty << 10) | ((uint32_t)tz << 20));
			}
	FS3_CUDA(cudaMalloc((void**)&s->ls_order, sizeof(uint32_t) * nt));
	FS3_CUDA(cudaMalloc((void**)&s->ls_done, sizeof(uint32_t) * nt));
	FS3_CUDA(cudaMalloc((void**)&s->ls_ticket, sizeof(uint32_t)));
	FS3_CUDA(cudaMemcpyAsync(s->ls_order, ord.data(), sizeof(uint32_t) * nt, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaMemsetAsync(s->ls_done, 0, sizeof(uint32_t) * nt, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	s->ls_ntiles = nt;
	s->ls_epoch = 0;
	int per_sm = 0, sms = 0;
	FS3_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_ls_sweep_dataflow, LS_T * LS_T * LS_T, 0));
	FS3_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device));
	long long blocks = (long long)std::max(1, per_sm) * sms;
	s->ls_blocks = (int)std::min<long long>(blocks, 8LL * nt);
	return FS3_OK;
}
