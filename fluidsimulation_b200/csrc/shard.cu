// z-slab sharding of FluidSim3::Simulate over the GPUs of one box.
//
// One rank per GPU.  Every rank holds full-size grid arrays (global indexing) and owns the cells / face nodes with z in
// [z0, z1) and the particles whose cell (int)roundf(n*Z) (S3D.cpp:288) lies there.  All device memory of a rank is one
// arena exported through ONE CUDA-IPC handle, so rank a reaches array X of rank b at  peer_base[b] + (X - arena).
// Data moves by stores / copies into the peers' arrays over NVLink (no staging, no library collective):
//   * SOR: a multi-process handle solves in ONE cooperative launch with two-plane halos: the neighbours' boundary planes of one
//     colour are READ in place over NVLink once per iteration, and a courier CTA raises two flags per neighbour and iteration
//     (project.cu, k_sor_persistent).  A lockstep group (and slabs under 4 planes) launch every half sweep: the sweep kernel
//     stores its boundary plane into the neighbours' ghost planes and raises their epoch flag from its last block
//     (k_sor_half<.., SHARD>);
//   * particle migration: emigrants are packed per destination and copied to the tail of the destination's live range;
//   * ghost layers for the P2G gather: one cell layer of sorted particles + its cell_start rows is PULLED from each
//     neighbour and spliced in directly below / above the own range, so the gather kernel is unchanged;
//   * level-set sweeps: ONE dataflow over all ranks' tiles; a tile on a slab boundary stores its boundary plane of records, its
//     phi plane and its completion epoch into the neighbouring rank's arrays (levelset.cu, LsCtx);
//   * P2G result and projected field: the planes inside the neighbours' extrapolation window / velocity halo go to those
//     neighbours; whole slabs (projected field, phi, closest) follow on a copy stream behind a second flag, for the getters.
// Ordering is by epoch flags in device memory written with st.release.sys by the producer and polled by a one-warp
// kernel (or the sweep kernel's first warp) of the consumer.  A step is a fixed sequence of PHASES, each "wait, work,
// signal"; a multi-process handle runs them back to back on its stream, a lockstep group (all ranks in one process on one
// stream, for single-GPU testing) runs phase k of every rank before phase k+1, which satisfies every wait by stream
// order -- waits then never spin and a wait that is not already satisfied is reported as an error (flags[2]).
// Nothing runs replicated: the 8 Gauss-Seidel level-set sweeps are pipelined across the slabs tile by tile, and the layered
// extrapolation runs on a window of FS3_EX_DEPTH planes around the slab, which is exact for every node the step consumes (p2g.cu).
#include "fs3_internal.cuh"
#include <algorithm>

static inline size_t up512(size_t b) { return (b + 511) / 512 * 512; }

// Upper bound of everything fs3_create allocates for one rank (each allocation rounded up to 512 B).
size_t fs3i_arena_bytes(int n, long long ptot, long long mig_cap, int world) {
	const size_t N = (size_t)n, nC = N * N * N, nF = (N + 1) * N * N, nH = 8 * nC;
	size_t t = 0;
	t += 6 * up512(nF * 4) + up512(nC * 4);                                   // U V W oU oV oW phi
	t += 2 * up512((nH + 1) * 4) + up512(nH * 4);                             // cell_count, cell_start, cell_fill
	t += up512(((nH + 1) / 4096 + 2) * 8 + 16);                               // scan_block
	t += up512(nC * 16) + up512(nC * 4);                                      // ls_rec, closest
	t += 6 * up512(nF * 4) + 3 * up512(nF);                                   // cdv, orderv, exmask
	t += up512(3 * (3 * (3 * N + 8) + 2) * 4) + up512(16 * 2 * 4096) + 512;   // bucket, res, dev_flags
	t += 4 * up512((size_t)ptot * 16) + 2 * up512((size_t)ptot * 4);          // pos, vel (x2), cellid, tmp_id
	const size_t hp = ((N + 1) / 2 + 3) / 4 * 4, total = hp * N * (N + 2), nvec = total / 2;
	t += 2 * (3 * up512(total * 8) + up512(total));                           // p, b, diag, mask per colour
	t += up512((nvec + 1) * 4) + 2 * (up512(nvec * 4) + up512((nvec + 1) * 4)); // SOR lists
	const size_t nt = ((N + 7) / 8) * ((N + 7) / 8) * ((N + 3) / 4);
	t += up512(8 * nt * 4) + up512(nt * 4) + 512;                             // level-set item table, epochs, ticket
	t += up512((size_t)world * (world + 1) * 4) + up512((2 * (size_t)world + 1) * 4) + up512((size_t)mig_cap * 4) + 2 * up512((size_t)mig_cap * 16);
	t += up512(FS3_BAR_WORDS * 4) + FS3_SOR_DONE_BYTES;                       // bar, sor_done
	return t + (1u << 20);
}

int fs3i_alloc(fs3_sim* s, void** p, size_t bytes) {
	if (!s->arena) {
		FS3_CUDA(cudaMalloc(p, bytes ? bytes : 1));
		return FS3_OK;
	}
	size_t b = up512(bytes ? bytes : 1);
	if (s->arena_used + b > s->arena_bytes) { s->err = "device arena of the sharded handle is too small (fs3i_arena_bytes)"; return FS3_ERR_STATE; }
	*p = s->arena + s->arena_used;
	s->arena_used += b;
	return FS3_OK;
}
void fs3i_free(fs3_sim* s, void* p) {
	if (p && !s->arena) cudaFree(p);
}

int fs3i_shard_alloc(fs3_sim* s) {
	int rc;
	const int w = s->world;
	if ((rc = fs3i_alloc(s, (void**)&s->mig_matrix, (size_t)w * (w + 1) * 4)) || (rc = fs3i_alloc(s, (void**)&s->mig_ctl, (2 * (size_t)w + 1) * 4)) ||
	    (rc = fs3i_alloc(s, (void**)&s->leaver_idx, (size_t)s->mig_cap * 4)) || (rc = fs3i_alloc(s, (void**)&s->mig_pos, (size_t)s->mig_cap * 16)) ||
	    (rc = fs3i_alloc(s, (void**)&s->mig_vel, (size_t)s->mig_cap * 16)) || (rc = fs3i_alloc(s, (void**)&s->bar, FS3_BAR_WORDS * 4)) ||
	    (rc = fs3i_alloc(s, (void**)&s->sor_done, FS3_SOR_DONE_BYTES)))   // [0] block counter of the end-of-sweep signal; from [32] on the one-launch solve's grid barrier
		return rc;
	FS3_CUDA(cudaMallocHost((void**)&s->h_pin, (size_t)std::max(w * (w + 1), 64) * 4));
	return FS3_OK;
}

// ---- barrier = signal (publish my epoch to every rank) + wait (every rank has published it) -----------------------
struct PeerPtrs { uint32_t* p[FS3_MAX_WORLD]; };

__global__ void k_bar_signal(PeerPtrs bars, int rank, int world, uint32_t epoch) {
	const int r = threadIdx.x;
	if (r < world && r != rank) { __threadfence_system(); fs3_st_release_sys(bars.p[r] + rank, epoch); }
}
__global__ void k_bar_wait(const uint32_t* bar, int rank, int world, uint32_t epoch, int* flags, int nospin) {
	const int r = threadIdx.x;
	if (r < world && r != rank) fs3_flag_wait(bar + r, epoch, flags, nospin);
}
// st: the stream the publication is ordered on (default: the handle's main stream)
int fs3i_signal(fs3_sim* s, cudaStream_t st) {
	PeerPtrs pp;
	for (int r = 0; r < s->world; r++) pp.p[r] = fs3_peer(s, r, s->bar);
	s->bar_epoch++;
	k_bar_signal<<<1, 32, 0, st ? st : s->stream>>>(pp, s->rank, s->world, s->bar_epoch); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_wait(fs3_sim* s, cudaStream_t st) {
	k_bar_wait<<<1, 32, 0, st ? st : s->stream>>>(s->bar, s->rank, s->world, s->bar_epoch, s->dev_flags, s->lockstep ? 1 : 0); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// ---- selective epoch flags: rank r's word of flag set `slot0` on the ranks in `mask` ---------------------------------------
__global__ void k_flag_signal(PeerPtrs bars, int rank, unsigned mask, int slot0, uint32_t epoch) {
	const int r = threadIdx.x;
	if (r < FS3_MAX_WORLD && ((mask >> r) & 1u) && r != rank) { __threadfence_system(); fs3_st_release_sys(bars.p[r] + slot0 + rank, epoch); }
}
__global__ void k_flag_wait(const uint32_t* bar, int rank, unsigned mask, int slot0, uint32_t epoch, int* flags, int nospin) {
	const int r = threadIdx.x;
	if (r < FS3_MAX_WORLD && ((mask >> r) & 1u) && r != rank) fs3_flag_wait(bar + slot0 + r, epoch, flags, nospin);
}
int fs3i_flag_signal(fs3_sim* s, unsigned mask, int slot0, uint32_t epoch, cudaStream_t st) {
	PeerPtrs pp;
	for (int r = 0; r < FS3_MAX_WORLD; r++) pp.p[r] = r < s->world ? fs3_peer(s, r, s->bar) : nullptr;
	k_flag_signal<<<1, 32, 0, st ? st : s->stream>>>(pp, s->rank, mask, slot0, epoch); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_flag_wait(fs3_sim* s, unsigned mask, int slot0, uint32_t epoch, cudaStream_t st) {
	k_flag_wait<<<1, 32, 0, st ? st : s->stream>>>(s->bar, s->rank, mask, slot0, epoch, s->dev_flags, s->lockstep ? 1 : 0); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
// The projected field of the last step is complete on this rank once every other rank's copy of its slab has landed (the
// copies run on the senders' copy streams behind the step): what the whole-grid getters wait for.
// My planes [a, b) of a z-slowest array go to every rank in `mask` whose range of needed planes [lo(r), hi(r)) overlaps them.
template <typename F> static int push_overlap(fs3_sim* s, void* arr, size_t plane_bytes, int a, int b, unsigned mask, F need, cudaStream_t st) {
	for (int k = 1; k < s->world; k++) {
		const int r = (s->rank + k) % s->world;   // staggered so that the ranks do not all target the same peer at once
		if (!((mask >> r) & 1u)) continue;
		int lo, hi;
		need(r, &lo, &hi);
		lo = std::max(lo, a); hi = std::min(hi, b);
		if (lo >= hi) continue;
		FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, r, (char*)arr) + plane_bytes * (size_t)lo, (char*)arr + plane_bytes * (size_t)lo, plane_bytes * (size_t)(hi - lo), cudaMemcpyDefault,
		                         st ? st : s->stream));
	}
	return FS3_OK;
}
// this rank's slab of the projected field, phi and closest into every rank's arrays, then the "whole slab pushed" flag
int fs3i_field_push(fs3_sim* s, cudaStream_t st) {
	const size_t pU = (size_t)(s->nx + 1) * s->ny, pV = (size_t)s->nx * (s->ny + 1), pW = (size_t)s->nx * s->ny, pC = pW;
	const unsigned all = (1u << s->world) - 1u;
	const int nz = s->nz, w_a = s->rank == 0 ? 0 : s->z0 + 1, w_b = s->z1 + 1;
	auto need_all_c = [=](int, int* lo, int* hi) { *lo = 0; *hi = nz; };
	auto need_all_w = [=](int, int* lo, int* hi) { *lo = 0; *hi = nz + 1; };
	int rc;
	s->gather_epoch++;
	if ((rc = push_overlap(s, s->U, pU * 4, s->z0, s->z1, all, need_all_c, st)) || (rc = push_overlap(s, s->V, pV * 4, s->z0, s->z1, all, need_all_c, st)) ||
	    (rc = push_overlap(s, s->W, pW * 4, w_a, w_b, all, need_all_w, st)) || (rc = push_overlap(s, s->phi, pC * 4, s->z0, s->z1, all, need_all_c, st)) ||
	    (rc = push_overlap(s, s->closest, pC * 4, s->z0, s->z1, all, need_all_c, st)) || (rc = fs3i_flag_signal(s, all, FS3_BAR_XALL, s->gather_epoch, st)))
		return rc;
	s->field_dirty = false;
	return FS3_OK;
}
// Whole-grid getters of a sharded handle (COLLECTIVE: every rank calls them in the same order): each rank pushes its slab to
// all the others if a step has run since the last gather, then waits until every other rank's slab has landed.
int fs3i_field_wait(fs3_sim* s) {
	if (s->world < 2) return FS3_OK;
	int rc;
	if (s->field_dirty && (rc = fs3i_field_push(s, s->stream))) return rc;
	if (!s->gather_epoch) return FS3_OK;
	return fs3i_flag_wait(s, (1u << s->world) - 1u, FS3_BAR_XALL, s->gather_epoch, nullptr);
}

// ranks (other than this one) whose slab lies within `halo` planes of this rank's slab
static unsigned near_mask(const fs3_sim* s, int halo) {
	const int slab = s->nz / s->world;
	unsigned m = 0;
	for (int r = 0; r < s->world; r++) {
		if (r == s->rank) continue;
		const int a = r * slab, b = a + slab;
		if (a < s->z1 + halo + 1 && b > s->z0 - halo - 1) m |= 1u << r;
	}
	return m;
}
// planes of the velocity halo: a rank reads the projected field up to n / 8 planes (at most one neighbouring slab, at least 8
// planes) away from its own: 16.7 m/s at dt = 0.01 whatever the grid.  Advect checks that no Runge-Kutta stage leaves that range.
int fs3i_velocity_halo(const fs3_sim* s) { return std::max(8, std::min(s->nz / s->world, s->nz / 8)); }

int fs3i_push(fs3_sim* s, void* ptr, size_t byte_off, size_t bytes, cudaStream_t st) {
	if (!bytes) return FS3_OK;
	for (int k = 1; k < s->world; k++) {
		const int r = (s->rank + k) % s->world;   // staggered so that the ranks do not all target the same peer at once
		FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, r, (char*)ptr) + byte_off, (char*)ptr + byte_off, bytes, cudaMemcpyDefault, st ? st : s->stream));
	}
	return FS3_OK;
}

// Slab all-gathers that the next kernels of this rank do not depend on go to stream3, behind everything queued on the main
// stream so far: the copies (and the barrier signal that follows them) overlap this rank's next kernels.  A lockstep group
// keeps them on the shared stream, where stream order is what satisfies the other ranks' waits.
static int xfer_begin(fs3_sim* s, cudaStream_t* st) {
	if (s->lockstep) { *st = s->stream; return FS3_OK; }
	FS3_CUDA(cudaEventRecord(s->ev_fork, s->stream));
	FS3_CUDA(cudaStreamWaitEvent(s->stream3, s->ev_fork, 0));
	*st = s->stream3;
	return FS3_OK;
}
// main stream waits until everything queued on stream3 so far has completed (before it overwrites what was being sent)
static int xfer_join(fs3_sim* s) {
	if (s->lockstep) return FS3_OK;
	FS3_CUDA(cudaEventRecord(s->ev_xfer, s->stream3));
	FS3_CUDA(cudaStreamWaitEvent(s->stream, s->ev_xfer, 0));
	return FS3_OK;
}

// ---- SOR neighbour epochs: slot FS3_MAX_WORLD+0 is written by the lower, +1 by the upper neighbour ----------------
__global__ void k_sor_signal(uint32_t* lo_flag, uint32_t* hi_flag, uint32_t epoch) {
	__threadfence_system();
	if (threadIdx.x == 0 && lo_flag) fs3_st_release_sys(lo_flag, epoch);
	if (threadIdx.x == 1 && hi_flag) fs3_st_release_sys(hi_flag, epoch);
}
__global__ void k_sor_wait(const uint32_t* my_flag, int has_lo, int has_hi, uint32_t epoch, int* flags, int nospin) {
	if (threadIdx.x == 0 && has_lo) fs3_flag_wait(my_flag + 0, epoch, flags, nospin);
	if (threadIdx.x == 1 && has_hi) fs3_flag_wait(my_flag + 1, epoch, flags, nospin);
}
int fs3i_sor_signal(fs3_sim* s) {
	s->sor_epoch++;
	s->sor_e0 = s->sor_epoch;
	uint32_t* lo = s->rank > 0 ? fs3_peer(s, s->rank - 1, s->bar) + FS3_MAX_WORLD + 1 : nullptr;
	uint32_t* hi = s->rank < s->world - 1 ? fs3_peer(s, s->rank + 1, s->bar) + FS3_MAX_WORLD + 0 : nullptr;
	k_sor_signal<<<1, 32, 0, s->stream>>>(lo, hi, s->sor_e0); FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_sor_wait(fs3_sim* s, uint32_t epoch) {
	k_sor_wait<<<1, 32, 0, s->stream>>>(s->bar + FS3_MAX_WORLD, s->rank > 0, s->rank < s->world - 1, epoch, s->dev_flags, s->lockstep ? 1 : 0);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// ---- particle migration ----------------------------------------------------------------------------------------------
// row `rank` of the count matrix -- emigrants per destination, then the length of my live range -- goes into every rank's copy
__global__ void k_mig_publish(PeerPtrs mats, const uint32_t* __restrict__ ctl, int rank, int world, uint32_t live_len) {
	const int d = threadIdx.x;
	if (d >= world) return;
	uint32_t* row = mats.p[d] + rank * (world + 1);
	for (int j = 0; j < world; j++) row[j] = ctl[j];
	row[world] = live_len;
}
int fs3i_mig_publish(fs3_sim* s) {
	PeerPtrs pp;
	for (int r = 0; r < s->world; r++) pp.p[r] = fs3_peer(s, r, s->mig_matrix);
	k_mig_publish<<<1, 32, 0, s->stream>>>(pp, s->mig_ctl, s->rank, s->world, (uint32_t)s->np_in); FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// After the barrier every rank holds the whole count matrix: bring it to the host (the one host round trip of the
// migration), append my emigrants to the tail of each destination's live range, and advance the replicated bookkeeping.
int fs3i_mig_exchange(fs3_sim* s) {
	const int w = s->world, me = s->rank, W1 = w + 1;
	FS3_CUDA(cudaMemcpyAsync(s->h_pin, s->mig_matrix, (size_t)w * W1 * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	int rc = fs3i_check_flags(s);
	if (rc) return rc;
	const uint32_t* m = s->h_pin;   // m[r*W1 + d] = particles r -> d;  m[r*W1 + w] = length of r's live range (hole-free before this migration)
	long long out_me = 0;
	for (int d = 0; d < w; d++) out_me += m[me * W1 + d];
	if (out_me > s->mig_cap) { s->err = "more emigrants in one step than the migration buffer holds"; return FS3_ERR_STATE; }
	long long base = 0;
	for (int d = 0; d < w; d++) {
		const long long cnt = m[me * W1 + d];
		if (d != me && cnt > 0) {
			long long off = m[d * W1 + w];
			for (int r = 0; r < me; r++) off += m[r * W1 + d];
			FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, d, s->pos[s->cur]) + s->poff + off, s->mig_pos + base, (size_t)cnt * sizeof(float4), cudaMemcpyDefault, s->stream));
			FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, d, s->vel[s->cur]) + s->poff + off, s->mig_vel + base, (size_t)cnt * sizeof(float4), cudaMemcpyDefault, s->stream));
		}
		base += cnt;
	}
	for (int d = 0; d < w; d++) {
		long long arrive = 0, leave = 0;
		for (int r = 0; r < w; r++) { arrive += m[r * W1 + d]; leave += m[d * W1 + r]; }
		const long long len = m[d * W1 + w];
		if (len + arrive > s->cap) { s->err = "particle capacity of a slab exceeded by migration"; return FS3_ERR_STATE; }
		if (d == me) s->np_in = len + arrive;
		s->np_all[d] = len - leave + arrive;
	}
	s->np = s->np_all[me];
	s->sorted = false; s->counted = false;
	return FS3_OK;
}

// ---- ghost layers: one cell layer (two half-cell layers) of each neighbour's sorted particles ---------------------
// Keys [ka, kb) of the neighbour's cell_start describe the layer.  Its particles are copied to just below (hi == 0) or just
// above (hi == 1) my own sorted range and my cell_start rows for those keys are rebased, so that cell_start stays one
// monotone array over [ka_lo, kb_hi] and every gather kernel works unchanged on slab + ghosts.
__global__ void __launch_bounds__(256) k_ghost_pull(const uint32_t* __restrict__ nstart, const float4* __restrict__ npos, const float4* __restrict__ nvel,
                                                    uint32_t* __restrict__ mystart, float4* __restrict__ mypos, float4* __restrict__ myvel, size_t ka,
                                                    size_t kb, int hi, uint32_t poff, uint32_t np, uint32_t gcap, int* __restrict__ flags) {
	const uint32_t nb = nstart[ka], ne = nstart[kb], cnt = ne - nb;
	if (cnt > gcap) { if (blockIdx.x == 0 && threadIdx.x == 0) flags[3] = 1; return; }
	const uint32_t dst0 = hi ? poff + np : poff - cnt;
	const size_t nk = kb - ka + (hi ? 1 : 0);
	const size_t stride = (size_t)gridDim.x * blockDim.x, t0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	for (size_t k = t0; k < nk; k += stride) mystart[ka + k] = dst0 + (nstart[ka + k] - nb);
	for (size_t t = t0; t < cnt; t += stride) { mypos[dst0 + t] = npos[nb + t]; myvel[dst0 + t] = nvel[nb + t]; }
}
int fs3i_ghost_pull(fs3_sim* s) {
	const size_t hplane = (size_t)(2 * s->nx) * (2 * s->ny);
	const int nb = s->sm_count * 4;
	for (int hi = 0; hi < 2; hi++) {
		const int r = hi ? s->rank + 1 : s->rank - 1;
		if (r < 0 || r >= s->world) continue;
		const size_t ka = hi ? hplane * (size_t)(2 * s->z1) : hplane * (size_t)(2 * s->z0 - 2), kb = ka + 2 * hplane;
		k_ghost_pull<<<nb, 256, 0, s->stream>>>(fs3_peer(s, r, s->cell_start), fs3_peer(s, r, s->pos[s->cur]), fs3_peer(s, r, s->vel[s->cur]), s->cell_start,
		                                         s->pos[s->cur], s->vel[s->cur], ka, kb, hi, (uint32_t)s->poff, (uint32_t)s->np, (uint32_t)s->gcap, s->dev_flags);
		FS3_LAUNCH_CHECK();
	}
	return FS3_OK;
}

// ---- the step as phases ------------------------------------------------------------------------------------------------
int fs3i_num_step_phases(const fs3_sim* s) { return 8 + 2 * s->prm.sor_iterations; }

static float flip_alpha(const fs3_sim* s, float dt) {
	float a = 6 * dt * s->prm.nu * s->cpm * s->cpm;   // S3D.cpp:114
	return a < 0.0f ? 0.0f : (a > 1.0f ? 1.0f : a);
}

// slab of a z-slowest array: element range [plane*za, plane*zb)
static int push_planes(fs3_sim* s, void* arr, size_t plane_elems, size_t elem_bytes, int za, int zb, cudaStream_t st = nullptr) {
	return fs3i_push(s, arr, plane_elems * elem_bytes * (size_t)za, plane_elems * elem_bytes * (size_t)(zb - za), st);
}

int fs3i_step_phase(fs3_sim* s, int ph, float dt) {
	const bool T = s->prm.record_stage_times != 0;
	const int I = s->prm.sor_iterations;
	const size_t pU = (size_t)(s->nx + 1) * s->ny, pV = (size_t)s->nx * (s->ny + 1), pW = (size_t)s->nx * s->ny, pC = pW;
	const bool last = s->rank == s->world - 1;
	int rc;
	if (ph == 0) {          // Advect (S3D.cpp:190-232); emigrants -> outgoing buffers; publish the count row
		if (T) cudaEventRecord(s->ev[FS3_T_ADVECT], s->stream);
		if ((rc = fs3i_advect(s, dt)) || (rc = fs3i_mig_pack(s)) || (rc = fs3i_mig_publish(s))) return rc;
		return fs3i_signal(s);
	}
	if (ph == 1) {          // counts of all ranks are here: send the emigrants
		if ((rc = fs3i_wait(s)) || (rc = fs3i_mig_exchange(s))) return rc;
		return fs3i_signal(s);
	}
	if (ph == 2) {          // immigrants have landed: counting sort of the slab's particles
		if ((rc = fs3i_wait(s))) return rc;
		if (T) cudaEventRecord(s->ev[FS3_T_BIN], s->stream);
		if ((rc = fs3i_bin(s))) return rc;
		return fs3i_signal(s);
	}
	if (ph == 3) {          // neighbours are sorted: ghost layers; level-set seed of the slab
		if ((rc = fs3i_wait(s)) || (rc = fs3i_ghost_pull(s)) || (rc = fs3i_levelset_seed(s))) return rc;
		// the copies of the previous step's phi / closest slabs (stream3) have left before the sweeps rewrite them
		if (!s->lockstep) { FS3_CUDA(cudaEventRecord(s->ev_xfer, s->stream3)); FS3_CUDA(cudaStreamWaitEvent(s->stream2, s->ev_xfer, 0)); }
		return FS3_OK;
	}
	if (ph == 4) {          // the 8 sweeps of the slab's tiles start on stream2 (they exchange boundary planes with the neighbours' sweep
		                    // kernels while they run); meanwhile P2G of the slab's face nodes on the main stream
		if ((rc = fs3i_levelset_fork(s, false))) return rc;
		// no wait on the main stream: the gather reads this rank's particles + ghost layers only.  Each face grid leaves as soon as
		// it is done (while the next one is gathered); invalid nodes travel as tagged NaNs, so the validity needs no copy of its own.
		// Only the ranks whose extrapolation window (FS3_EX_DEPTH planes around their slab) reaches into this slab get planes.
		const int zw = s->z1 + (last ? 1 : 0);
		const int slab = s->nz / s->world, world = s->world, nz = s->nz;
		const unsigned near = near_mask(s, FS3_EX_DEPTH);
		cudaStream_t xs = nullptr;
		for (int ax = 0; ax < 3; ax++) {
			if ((rc = fs3i_p2g_gather_axis(s, ax)) || (rc = xfer_begin(s, &xs))) return rc;
			auto need = [=](int r, int* lo, int* hi) {
				const int top = nz + (ax == 2 ? 1 : 0), own1 = (r + 1) * slab + ((ax == 2 && r == world - 1) ? 1 : 0);
				*lo = std::max(0, r * slab - FS3_EX_DEPTH); *hi = std::min(top, own1 + FS3_EX_DEPTH);
			};
			if (ax == 0) rc = push_overlap(s, s->oU, pU * 4, s->z0, s->z1, near, need, xs);
			else if (ax == 1) rc = push_overlap(s, s->oV, pV * 4, s->z0, s->z1, near, need, xs);
			else rc = push_overlap(s, s->oW, pW * 4, s->z0, zw, near, need, xs);
			if (rc) return rc;
		}
		return fs3i_signal(s, xs);
	}
	if (ph == 5) {          // extrapolation of the slab window beside the sweeps; pressure system of the slab; tell the neighbours it is set up
		// own outgoing copies are complete before the extrapolation overwrites the arrays they read
		if ((rc = xfer_join(s)) || (rc = fs3i_wait(s)) || (rc = fs3i_cd_from_tags(s)) || (rc = fs3i_extrapolate_all(s)) || (rc = fs3i_levelset_join(s)) ||
		    (rc = fs3i_project_setup(s, dt, true, true)))
			return rc;
		if (fs3i_sor_two_plane(s)) {
			// the red list also covers the neighbours' nearest plane: their masks must be here before it is built
			if ((rc = fs3i_sor_push_system(s)) || (rc = fs3i_sor_signal(s)) || (rc = fs3i_sor_wait(s, s->sor_e0)) || (rc = fs3i_sor_prepare(s))) return rc;
		} else if ((rc = fs3i_sor_prepare(s)) || (rc = fs3i_sor_signal(s))) return rc;
		if (T) cudaEventRecord(s->ev[FS3_T_PROJECT_SOR], s->stream);
		return FS3_OK;
	}
	if (ph < 6 + 2 * I) {   // compute + halo exchange in one kernel
		s->sor_halves = 2 * I;
		// one launch per half sweep for a lockstep group (the ranks' kernels take turns on one stream) and for slabs too thin for
		// two-plane halos; a multi-process handle runs the whole solve as ONE cooperative launch
		if (s->lockstep || !fs3i_sor_two_plane(s)) return fs3i_sor_half_launch(s, ph - 6);
		return ph == 6 ? fs3i_sor_persistent(s, I) : FS3_OK;
	}
	if (ph == 6 + 2 * I) {  // pressure gradient on the slab; the projected field is gathered on every rank
		// the copies of the previous step's field (stream3) have left before the gradient overwrites their source
		if ((rc = xfer_join(s)) || (!fs3i_sor_two_plane(s) && (rc = fs3i_sor_wait(s, s->sor_epoch))) || (rc = fs3i_project_gradient(s, dt, true, true))) return rc;
		// The gradient kernel of cell plane z writes the W faces z+1 (and z = 0 on rank 0): that is the range it owns here.
		// On the copy stream: the planes inside the neighbours' velocity halo + their flag -- all the next kernels of any rank
		// read (FLIP blend, G2P, the next Advect).  Whole slabs only travel when a whole-grid getter asks (fs3i_field_wait).
		const int hv = fs3i_velocity_halo(s), slab = s->nz / s->world, nz = s->nz;
		const unsigned near = near_mask(s, hv);
		const int w_a = s->rank == 0 ? 0 : s->z0 + 1, w_b = s->z1 + 1;
		auto need_c = [=](int r, int* lo, int* hi) { *lo = std::max(0, r * slab - hv); *hi = std::min(nz, (r + 1) * slab + hv); };
		auto need_w = [=](int r, int* lo, int* hi) { *lo = std::max(0, r * slab - hv); *hi = std::min(nz + 1, (r + 1) * slab + hv + 1); };
		cudaStream_t xs;
		if ((rc = xfer_begin(s, &xs))) return rc;
		s->xfer_epoch++;
		if ((rc = push_overlap(s, s->U, pU * 4, s->z0, s->z1, near, need_c, xs)) || (rc = push_overlap(s, s->V, pV * 4, s->z0, s->z1, near, need_c, xs)) ||
		    (rc = push_overlap(s, s->W, pW * 4, w_a, w_b, near, need_w, xs)) || (rc = fs3i_flag_signal(s, near, FS3_BAR_XNEAR, s->xfer_epoch, xs)))
			return rc;
		s->field_dirty = true;
		// a lockstep group gathers right away (its getters are called rank by rank, so they cannot wait for each other); a
		// multi-process rank gathers when a whole-grid getter is called (fs3i_field_wait): nothing in a step reads those planes
		if (s->lockstep) return fs3i_field_push(s, xs);
		return FS3_OK;
	}
	if (ph == 7 + 2 * I) {  // PIC/FLIP blend (S3D.cpp:148-180), once the neighbours' halo planes of the projected field are here
		if ((rc = fs3i_flag_wait(s, near_mask(s, fs3i_velocity_halo(s)), FS3_BAR_XNEAR, s->xfer_epoch))) return rc;
		if (T) cudaEventRecord(s->ev[FS3_T_FLIP_APPLY], s->stream);
		const float alpha = flip_alpha(s, dt);
		if ((rc = fs3i_flip_combine(s, alpha)) || (rc = fs3i_g2p(s, alpha))) return rc;
		if (T) cudaEventRecord(s->ev[FS3_T_TOTAL], s->stream);
		return FS3_OK;
	}
	s->err = "no such step phase";
	return FS3_ERR_INVALID;
}

// ResetSimulation (S3D.cpp:41-99) as phases: generate the slab's particles, migrate the ones whose jitter rounded them
// into a neighbouring slab, sort.
int fs3i_reset_phase(fs3_sim* s, int ph) {
	int rc;
	if (ph == 0) {
		if ((rc = fs3i_reset_grids(s)) || (rc = fs3i_reset_particles_slab(s)) || (rc = fs3i_mig_mark(s)) || (rc = fs3i_mig_pack(s)) ||
		    (rc = fs3i_mig_publish(s)))
			return rc;
		return fs3i_signal(s);
	}
	if (ph == 1) {
		if ((rc = fs3i_wait(s)) || (rc = fs3i_mig_exchange(s))) return rc;
		return fs3i_signal(s);
	}
	if (ph == 2) {
		if ((rc = fs3i_wait(s))) return rc;
		return fs3i_bin(s);
	}
	s->err = "no such reset phase";
	return FS3_ERR_INVALID;
}
