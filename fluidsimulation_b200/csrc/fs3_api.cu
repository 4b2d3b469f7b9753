// C ABI of libfluidsim3_b200.so (include/fs3.h): lifecycle, state access, step orchestration.
// Mirrors FluidSim3's surface (Simulation3D.h:18-24,160-162); no CPU fallback anywhere.
#include "fs3_internal.cuh"
#include <cstring>
#include <cmath>
#include <new>
#include <algorithm>

static std::string g_create_error;
void fs3i_set_global_error(const std::string& msg) { g_create_error = msg; }   // host-only entry points have no handle to put it in

#define ENTER(s)                                                       \
	if (!(s)) return FS3_ERR_INVALID;                                  \
	do {                                                               \
		cudaError_t e0_ = cudaSetDevice((s)->device);                  \
		if (e0_ != cudaSuccess) { (s)->err = cudaGetErrorString(e0_); return FS3_ERR_CUDA; } \
	} while (0)

static int fail(fs3_sim* s, int code, const char* msg) { s->err = msg; return code; }

static int ensure_particle_capacity(fs3_sim* s, long long count) {
	if ((unsigned long long)count > (unsigned long long)s->nH) { s->err = "more particles than half-cells (8 per cell on average): not supported"; return FS3_ERR_INVALID; }
	if (count <= s->cap) return FS3_OK;
	if (s->arena) { s->err = "particle capacity of a sharded handle is fixed at creation"; return FS3_ERR_STATE; }
	for (int k = 0; k < 2; k++) {
		if (s->pos[k]) cudaFree(s->pos[k]); if (s->vel[k]) cudaFree(s->vel[k]);
		s->pos[k] = s->vel[k] = nullptr;
	}
	if (s->cellid) cudaFree(s->cellid); if (s->tmp_id) cudaFree(s->tmp_id);
	s->cellid = nullptr; s->tmp_id = nullptr;
	size_t n = (size_t)count;
	for (int k = 0; k < 2; k++) { FS3_CUDA(cudaMalloc((void**)&s->pos[k], n * sizeof(float4))); FS3_CUDA(cudaMalloc((void**)&s->vel[k], n * sizeof(float4))); }
	FS3_CUDA(cudaMalloc((void**)&s->cellid, n * sizeof(uint32_t)));
	FS3_CUDA(cudaMalloc((void**)&s->tmp_id, n * sizeof(int)));
	s->cap = count;
	return FS3_OK;
}

// device scratch for AoS staging at the boundary, owned by the handle
int fs3i_staging(fs3_sim* s, size_t bytes) {
	if (bytes <= s->stage_bytes) return FS3_OK;
	if (s->stage_ptr) cudaFree(s->stage_ptr);
	s->stage_ptr = nullptr; s->stage_bytes = 0;
	FS3_CUDA(cudaMalloc(&s->stage_ptr, bytes));
	s->stage_bytes = bytes;
	return FS3_OK;
}

int fs3i_check_flags(fs3_sim* s) {
	int h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
	FS3_CUDA(cudaMemcpyAsync(h, s->dev_flags, 8 * sizeof(int), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	if (h[0] || h[1] || h[2] || h[3] || h[4]) FS3_CUDA(cudaMemsetAsync(s->dev_flags, 0, 8 * sizeof(int), s->stream));
	if (h[4]) return fail(s, FS3_ERR_STATE, "a particle's Runge-Kutta stage left the velocity halo of its slab (flow faster than one slab per step); results invalid");
	if (h[2]) return fail(s, FS3_ERR_STATE, "a cross-rank wait was not satisfied (peer rank missing, out of step, or timed out); results invalid");
	if (h[3]) return fail(s, FS3_ERR_STATE, "a slab's particle, ghost-layer or migration capacity was exceeded; results invalid");
	if (h[1]) return fail(s, FS3_ERR_STATE, "a device-side dependency wait timed out (level-set tile hand-off or extrapolation layer barrier); results invalid");
	if (h[0]) return fail(s, FS3_ERR_PARTICLE, "a particle lay outside the grid (or its rank's slab) when binned (positions must be inside the Advect clamp box)");
	return FS3_OK;
}

extern "C" {

const char* fs3_last_error(const fs3_sim* s) { return s ? s->err.c_str() : g_create_error.c_str(); }

// The reference's hard-coded constants (S3D.h:135-145, S3D.cpp:104,944-946).
int fs3_default_params(fs3_params* p) {
	if (!p) return FS3_ERR_INVALID;
	memset((void*)p, 0, sizeof(*p));
	p->sor_iterations = 100; p->omega = 0; p->precision = 0; p->force_dt_0_01 = 1; p->record_residuals = 0;
	p->rho = 1000.0f; p->gravity_y = -9.81f; p->nu = 8.90f * 1e-4f / 1000.0f; p->particle_radius = 1.0f;
	p->record_stage_times = 0;
	return FS3_OK;
}

// Shared by fs3_create (world 1), fs3_create_sharded (one rank of a multi-process job) and fs3_group_create (all ranks in
// this process).  A sharded handle is NOT reset here: ResetSimulation needs the peers (fs3_shard_connect / group create).
static int create_common(int nx, int ny, int nz, float cpm, int device, int rank, int world, bool lockstep, cudaStream_t shared_stream, fs3_sim** out) {
	if (!out) return FS3_ERR_INVALID;
	*out = nullptr;
	if (nx < 4 || nx != ny || ny != nz || cpm != (float)nx) {
		g_create_error = "only cubic grids with cells_per_meter == nx (>= 4) are supported, as the reference's own results are only defined there";
		return FS3_ERR_INVALID;
	}
	if (nx > 812) {
		g_create_error = "grids above 812^3 are not supported: half-cell sort keys ((2n)^3) and grid element indices are 32 bits";
		return FS3_ERR_INVALID;
	}
	if (world < 1 || world > FS3_MAX_WORLD || rank < 0 || rank >= world || nz % world != 0 || (world > 1 && nz / world < 2)) {
		g_create_error = "z-slab sharding needs 1 <= world <= 16, nz divisible by world and slabs of at least 2 cell planes";
		return FS3_ERR_INVALID;
	}
	int ndev = 0;
	cudaError_t e = cudaGetDeviceCount(&ndev);
	if (e != cudaSuccess || device < 0 || device >= ndev) {
		g_create_error = std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range");
		return FS3_ERR_CUDA;
	}
	cudaDeviceProp prop;
	cudaGetDeviceProperties(&prop, device);
	if (prop.major != 10) {
		g_create_error = std::string("device is not sm_100-class (found ") + prop.name + "); this library contains sm_100a code only";
		return FS3_ERR_CUDA;
	}
	fs3_sim* s = new (std::nothrow) fs3_sim();
	if (!s) return FS3_ERR_INVALID;
	fs3_default_params(&s->prm);
	s->device = device; s->nx = nx; s->ny = ny; s->nz = nz; s->cpm = cpm;
	s->nU = (size_t)(nx + 1) * ny * nz; s->nV = (size_t)nx * (ny + 1) * nz; s->nW = (size_t)nx * ny * (nz + 1); s->nC = (size_t)nx * ny * nz;
	s->launches = 0; s->steps = 0; s->np = 0; s->np_in = 0; s->cap = 0; s->cur = 0; s->sorted = false; s->counted = false; s->closest_valid = false; s->system_valid = false;
	s->sor_list_valid = false; s->sm_count = prop.multiProcessorCount; s->sor_per_sm = 0; s->persist_cap = -1;
	s->alloc_precision = -1; s->res_cap = 4096; s->res_count = 0; s->stage_ptr = nullptr; s->stage_bytes = 0;
	s->rank = rank; s->world = world; s->z0 = rank * (nz / world); s->z1 = s->z0 + nz / world; s->lockstep = lockstep;
	s->bar_epoch = 0; s->sor_epoch = 0; s->xfer_epoch = 0; s->gather_epoch = 0; s->field_dirty = false; s->sor_e0 = 0; s->n_ipc_opened = 0; s->poff = 0; s->gcap = 0; s->mig_cap = 0;
	for (int k = 0; k < FS3_T_COUNT; k++) s->stage_ms[k] = 0.f;
	const long long P = std::max(1LL, 8LL * (nz - 2) * (ny - 2) * (nx - 1 - nx / 2));
	long long cap = P;
	if (world > 1) {
		// a slab may hold up to twice its even share; a ghost region up to 8 average cell layers; a step's emigrants up to 4
		const long long layer = P / (nz - 2) + 1;
		cap = std::min(P, 2 * (P / world) + 4096);
		s->gcap = std::min(P, 8 * layer + 4096);
		s->mig_cap = std::min(P, 4 * layer + 4096);
		s->poff = (size_t)s->gcap;
	}
	const long long ptot = cap + 2 * s->gcap;
#define CK(call)                                                                                  \
	do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { g_create_error = std::string(#call) + ": " + cudaGetErrorString(e_); fs3_destroy(s); return FS3_ERR_CUDA; } } while (0)
#define CA(ptr, bytes)                                                                            \
	do { int rc_ = fs3i_alloc(s, (void**)&(ptr), (bytes)); if (rc_) { g_create_error = s->err; fs3_destroy(s); return rc_; } } while (0)
	CK(cudaSetDevice(device));
	if (shared_stream) { s->stream = shared_stream; s->owns_stream = false; }
	else { CK(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)); s->owns_stream = true; }
	for (int k = 0; k < FS3_T_COUNT + 2; k++) CK(cudaEventCreate(&s->ev[k]));
	CK(cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking));
	CK(cudaStreamCreateWithFlags(&s->stream3, cudaStreamNonBlocking));
	CK(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming)); CK(cudaEventCreate(&s->ev_ls_end)); CK(cudaEventCreateWithFlags(&s->ev_xfer, cudaEventDisableTiming));
	if (world > 1) {
		s->arena_bytes = fs3i_arena_bytes(nx, ptot, s->mig_cap, world);
		CK(cudaMalloc((void**)&s->arena, s->arena_bytes));
		CK(cudaMemsetAsync(s->arena, 0, s->arena_bytes, s->stream));   // epoch flags, counters and ghost planes start at zero
		s->arena_used = 0;
		s->peer_base[rank] = s->arena;
	}
	size_t nmax = s->nW > s->nU ? s->nW : s->nU;
	if (s->nV > nmax) nmax = s->nV;
	CA(s->U, s->nU * 4); CA(s->V, s->nV * 4); CA(s->W, s->nW * 4);
	CA(s->oU, s->nU * 4); CA(s->oV, s->nV * 4); CA(s->oW, s->nW * 4);
	CA(s->phi, s->nC * 4);
	CK(cudaMemsetAsync(s->phi, 0, s->nC * 4, s->stream));  // the reference leaves m_Phi uninitialised; every cell is written by the first level set
	s->nH = 8 * s->nC;
	CA(s->cell_count, (s->nH + 1) * 4); CA(s->cell_start, (s->nH + 1) * 4);
	CA(s->cell_fill, s->nH * 4);
	s->scan_block_cap = (s->nH + 1) / 4096 + 2;
	CA(s->scan_block, s->scan_block_cap * 8 + 16);   // 64-bit tile states of the look-back scan + its ticket
	CA(s->ls_rec, s->nC * sizeof(float4)); CA(s->closest, s->nC * 4);
	for (int k = 0; k < 3; k++) { CA(s->cdv[k], nmax * 4); CA(s->orderv[k], nmax * 4); CA(s->exmask[k], nmax); }
	s->bucket_cap = nx + ny + nz + 8;
	CA(s->bucket, 3 * (3 * (size_t)s->bucket_cap + 2) * 4);
	CA(s->ex_arrive, 4);
	CA(s->res, sizeof(double) * 2 * s->res_cap);
	CA(s->dev_flags, 8 * sizeof(int));
	CK(cudaMemsetAsync(s->dev_flags, 0, 8 * sizeof(int), s->stream));
	if (world > 1) {
		for (int k = 0; k < 2; k++) { CA(s->pos[k], (size_t)ptot * sizeof(float4)); CA(s->vel[k], (size_t)ptot * sizeof(float4)); }
		CA(s->cellid, (size_t)ptot * 4); CA(s->tmp_id, (size_t)ptot * 4);
		s->cap = cap;
	}
#undef CK
#undef CA
	int rc = world > 1 ? fs3i_shard_alloc(s) : ensure_particle_capacity(s, P);
	if (!rc) rc = fs3i_alloc_pressure(s);
	if (!rc) rc = fs3i_levelset_init(s);
	if (!rc && world == 1) rc = fs3i_reset(s);
	if (!rc && cudaStreamSynchronize(s->stream) != cudaSuccess) { s->err = "device initialisation failed"; rc = FS3_ERR_CUDA; }
	if (rc) { g_create_error = s->err; fs3_destroy(s); return rc; }
	*out = s;
	return FS3_OK;
}

int fs3_create(int nx, int ny, int nz, float cpm, int device, fs3_sim** out) { return create_common(nx, ny, nz, cpm, device, 0, 1, false, nullptr, out); }

// ---- z-slab sharding: one rank of a multi-process job -----------------------------------------------------------------
int fs3_create_sharded(int nx, int ny, int nz, float cpm, int device, int rank, int world, fs3_sim** out) {
	if (world == 1) return fs3_create(nx, ny, nz, cpm, device, out);
	return create_common(nx, ny, nz, cpm, device, rank, world, false, nullptr, out);
}

int fs3_shard_plan(int nz, int world, int rank, int* z0, int* z1) {
	if (world < 1 || world > FS3_MAX_WORLD || rank < 0 || rank >= world || nz % world != 0 || (world > 1 && nz / world < 2)) return FS3_ERR_INVALID;
	if (z0) *z0 = rank * (nz / world);
	if (z1) *z1 = (rank + 1) * (nz / world);
	return FS3_OK;
}
int fs3_shard_info(const fs3_sim* s, int* rank, int* world, int* z0, int* z1) {
	if (!s) return FS3_ERR_INVALID;
	if (rank) *rank = s->rank; if (world) *world = s->world; if (z0) *z0 = s->z0; if (z1) *z1 = s->z1;
	return FS3_OK;
}

// What a rank tells its peers: the CUDA-IPC handle of its arena plus enough identity to refuse unsafe set-ups.
struct fs3_shard_blob {
	cudaIpcMemHandle_t mem;       // 64 bytes
	unsigned char uuid[16];       // device the arena lives on
	unsigned long long arena_bytes;
	int rank, world, nx, pad;
};
static_assert(sizeof(fs3_shard_blob) <= FS3_SHARD_HANDLE_BYTES, "FS3_SHARD_HANDLE_BYTES too small");

int fs3_shard_get_handle(fs3_sim* s, void* blob) {
	ENTER(s);
	if (!blob) return FS3_ERR_INVALID;
	if (s->world < 2 || s->lockstep) return fail(s, FS3_ERR_STATE, "not a multi-process sharded handle");
	fs3_shard_blob b;
	memset(&b, 0, sizeof(b));
	FS3_CUDA(cudaIpcGetMemHandle(&b.mem, s->arena));
	cudaDeviceProp prop;
	FS3_CUDA(cudaGetDeviceProperties(&prop, s->device));
	memcpy(b.uuid, &prop.uuid, 16);
	b.arena_bytes = s->arena_bytes; b.rank = s->rank; b.world = s->world; b.nx = s->nx;
	memset(blob, 0, FS3_SHARD_HANDLE_BYTES);
	memcpy(blob, &b, sizeof(b));
	return FS3_OK;
}

static int run_reset_phases(fs3_sim* s) {
	for (int ph = 0; ph < 3; ph++) { int rc = fs3i_reset_phase(s, ph); if (rc) return rc; }
	return FS3_OK;
}

// `blobs` = world blobs of FS3_SHARD_HANDLE_BYTES each, in rank order (gathered by the caller with whatever transport
// it has: MPI_Allgather, torch.distributed.all_gather, a file).  Collective: maps every peer's arena, then runs
// ResetSimulation across the ranks.
int fs3_shard_connect(fs3_sim* s, const void* blobs) {
	ENTER(s);
	if (!blobs) return FS3_ERR_INVALID;
	if (s->world < 2 || s->lockstep) return fail(s, FS3_ERR_STATE, "not a multi-process sharded handle");
	if (s->n_ipc_opened) return fail(s, FS3_ERR_STATE, "already connected");
	fs3_shard_blob mine;
	memcpy(&mine, (const char*)blobs + (size_t)s->rank * FS3_SHARD_HANDLE_BYTES, sizeof(mine));
	for (int r = 0; r < s->world; r++) {
		fs3_shard_blob b;
		memcpy(&b, (const char*)blobs + (size_t)r * FS3_SHARD_HANDLE_BYTES, sizeof(b));
		if (b.rank != r || b.world != s->world || b.nx != s->nx || b.arena_bytes != s->arena_bytes)
			return fail(s, FS3_ERR_INVALID, "peer handle does not describe rank r of the same job (rank order, world size or grid differ)");
		if (r == s->rank) continue;
		// ranks spin on each other's flags inside kernels: two ranks on one GPU could wait on a kernel that cannot run
		if (!memcmp(b.uuid, mine.uuid, 16)) return fail(s, FS3_ERR_INVALID, "two ranks share one GPU; every rank needs its own device");
		void* base = nullptr;
		cudaError_t e = cudaIpcOpenMemHandle(&base, b.mem, cudaIpcMemLazyEnablePeerAccess);
		if (e != cudaSuccess) { s->err = std::string("cudaIpcOpenMemHandle (is peer access / IPC available between the GPUs?): ") + cudaGetErrorString(e); return FS3_ERR_CUDA; }
		s->peer_base[r] = (char*)base;
		s->n_ipc_opened++;
	}
	return run_reset_phases(s);
}

// ---- z-slab sharding: all ranks in THIS process, one stream, stepped in lockstep (single-GPU testing of the sharded path) --
int fs3_group_create(int nx, int ny, int nz, float cpm, int device, int world, fs3_sim** out) {
	if (!out || world < 2 || world > FS3_MAX_WORLD) return FS3_ERR_INVALID;
	for (int r = 0; r < world; r++) out[r] = nullptr;
	int rc = FS3_OK;
	for (int r = 0; r < world && !rc; r++) rc = create_common(nx, ny, nz, cpm, device, r, world, true, r ? out[0]->stream : nullptr, &out[r]);
	if (!rc) {
		for (int r = 0; r < world; r++) for (int q = 0; q < world; q++) out[r]->peer_base[q] = out[q]->arena;
		rc = fs3_group_reset(out, world);
		if (rc) g_create_error = "group reset failed";
	}
	if (rc) { for (int r = world - 1; r >= 0; r--) if (out[r]) { fs3_destroy(out[r]); out[r] = nullptr; } }
	return rc;
}
int fs3_group_reset(fs3_sim** h, int world) {
	if (!h || world < 2 || !h[0] || world != h[0]->world) return FS3_ERR_INVALID;
	cudaSetDevice(h[0]->device);
	for (int ph = 0; ph < 3; ph++)
		for (int r = 0; r < world; r++) { int rc = fs3i_reset_phase(h[r], ph); if (rc) return rc; }
	for (int r = 0; r < world; r++) h[r]->steps = 0;
	return FS3_OK;
}

int fs3_destroy(fs3_sim* s) {
	if (!s) return FS3_OK;
	cudaSetDevice(s->device);
	if (s->stream) cudaStreamSynchronize(s->stream);
	if (s->stream2) { cudaStreamSynchronize(s->stream2); cudaStreamDestroy(s->stream2); }
	if (s->stream3) { cudaStreamSynchronize(s->stream3); cudaStreamDestroy(s->stream3); }
	if (s->ev_fork) cudaEventDestroy(s->ev_fork); if (s->ev_join) cudaEventDestroy(s->ev_join); if (s->ev_ls_end) cudaEventDestroy(s->ev_ls_end); if (s->ev_xfer) cudaEventDestroy(s->ev_xfer);
	if (s->arena) {
		for (int r = 0; r < s->world; r++) if (!s->lockstep && r != s->rank && s->peer_base[r]) cudaIpcCloseMemHandle(s->peer_base[r]);
		cudaFree(s->arena);
		if (s->h_pin) cudaFreeHost(s->h_pin);
	} else {
		void* ptrs[] = {s->U, s->V, s->W, s->oU, s->oV, s->oW, s->phi, s->pos[0], s->pos[1], s->vel[0], s->vel[1], s->cellid, s->cell_count,
		                s->cell_start, s->cell_fill, s->tmp_id, s->scan_block, s->ls_rec, s->closest, s->cdv[0], s->cdv[1], s->cdv[2], s->orderv[0],
		                s->orderv[1], s->orderv[2], s->exmask[0], s->exmask[1], s->exmask[2], s->bucket, s->p[0], s->p[1], s->b[0], s->b[1], s->dg[0], s->dg[1], s->mask[0], s->mask[1],
		                s->res, s->dev_flags, s->ex_arrive, s->ls_order, s->ls_done, s->ls_ticket, s->sor_flags, s->sor_pos[0], s->sor_pos[1], s->sor_list[0], s->sor_list[1]};
		for (void* p : ptrs) if (p) cudaFree(p);
	}
	if (s->stage_ptr) cudaFree(s->stage_ptr);
	fs3i_consumer_free(s);
	for (int k = 0; k < FS3_T_COUNT + 2; k++) if (s->ev[k]) cudaEventDestroy(s->ev[k]);
	if (s->stream && s->owns_stream) cudaStreamDestroy(s->stream);
	delete s;
	return FS3_OK;
}

int fs3_reset(fs3_sim* s) {
	ENTER(s);
	if (s->world > 1) {
		if (s->lockstep) return fail(s, FS3_ERR_STATE, "a lockstep group is reset with fs3_group_reset");
		if (!s->n_ipc_opened) return fail(s, FS3_ERR_STATE, "sharded handle is not connected (fs3_shard_connect)");
		s->steps = 0;
		return run_reset_phases(s);   // collective: every rank calls fs3_reset
	}
	long long np = 8LL * (s->nz - 2) * (s->ny - 2) * (s->nx - 1 - s->nx / 2);
	int rc = ensure_particle_capacity(s, np > 0 ? np : 1);
	if (rc) return rc;
	s->steps = 0;
	return fs3i_reset(s);
}

int fs3_sync(fs3_sim* s) {
	ENTER(s);
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return fs3i_check_flags(s);
}

void* fs3_stream(fs3_sim* s) { return s ? (void*)s->stream : nullptr; }
long long fs3_kernel_launches(const fs3_sim* s) { return s ? s->launches : 0; }
long long fs3_num_particles(const fs3_sim* s) {
	if (!s) return 0;
	if (s->world == 1) return s->np;
	long long t = 0;
	for (int r = 0; r < s->world; r++) t += s->np_all[r];
	return t;   // m_particles.size() of the whole simulation
}
long long fs3_num_local_particles(const fs3_sim* s) { return s ? s->np : 0; }

static float flip_alpha(const fs3_sim* s, float dt) {
	float a = 6 * dt * s->prm.nu * s->cpm * s->cpm;   // S3D.cpp:114
	return a < 0.0f ? 0.0f : (a > 1.0f ? 1.0f : a);
}
static float effective_dt(const fs3_sim* s, float dt) {
	dt = dt < 0.0f ? 0.0f : (dt > 1.0f / 15.0f ? 1.0f / 15.0f : dt);   // S3D.cpp:103
	if (s->prm.force_dt_0_01) dt = 0.01f;                              // S3D.cpp:104
	return dt;
}

// FluidSim3::Simulate (S3D.cpp:101-188)
int fs3_step(fs3_sim* s, float dt) {
	ENTER(s);
	if (s->world == 1 && s->np <= 272) return fail(s, FS3_ERR_STATE, "Simulate needs more than 272 particles (the reference reads particles[272], Simulation3D.cpp:198-199)");
	dt = effective_dt(s, dt);
	const bool T = s->prm.record_stage_times != 0;
	int rc;
	if (s->world > 1) {
		if (s->lockstep) return fail(s, FS3_ERR_STATE, "a lockstep group is stepped with fs3_group_step");
		if (!s->n_ipc_opened) return fail(s, FS3_ERR_STATE, "sharded handle is not connected (fs3_shard_connect)");
		const int nph = fs3i_num_step_phases(s);
		for (int ph = 0; ph < nph; ph++) if ((rc = fs3i_step_phase(s, ph, dt))) return rc;
		s->steps++;
		return FS3_OK;
	}
	if (T) cudaEventRecord(s->ev[FS3_T_ADVECT], s->stream);
	if ((rc = fs3i_advect(s, dt))) return rc;
	if (T) cudaEventRecord(s->ev[FS3_T_BIN], s->stream);
	if ((rc = fs3i_bin(s))) return rc;
	s->ls_seed_forked = false;
	if (s->prm.levelset_mode == 1) {
		// line-sweep level set (opt-in, levelset_lines.cu): 26 short launches on the main stream, nothing to fork or join
		if ((rc = fs3i_levelset_lines(s)) || (rc = fs3i_p2g_gather(s)) || (rc = fs3i_extrapolate_all(s))) return rc;
	} else {
		if ((rc = fs3i_levelset_seed(s))) return rc;       // records LEVELSET_SEED
		// The 8 sweeps are bound by their dependency chain, not by throughput: they run on stream2 (LEVELSET_SWEEP is recorded
		// there) with ONE CTA per SM beside the P2G gather and the extrapolation, and are joined before the projection, which
		// is the first stage that reads phi.  At one CTA per SM the issue-bound gather loses 1.2 ms at 256^3 and the sweeps
		// take 6.6 instead of 3.7 ms, but all of that is hidden (19.55 -> 18.9 ms per step; 128^3: 4.24 -> 3.78; 512^3: even).
		if ((rc = fs3i_levelset_fork(s, false, false))) return rc;
		if ((rc = fs3i_p2g_gather(s))) return rc;          // records P2G
		if ((rc = fs3i_extrapolate_all(s))) return rc;
		if ((rc = fs3i_levelset_join(s))) return rc;
	}
	if ((rc = fs3i_project(s, dt, true, true))) return rc;  // records PROJECT_SETUP / SOR / TOVEL; gravity folded in
	if (T) cudaEventRecord(s->ev[FS3_T_FLIP_APPLY], s->stream);
	float alpha = flip_alpha(s, dt);
	if ((rc = fs3i_flip_combine(s, alpha))) return rc;
	if ((rc = fs3i_g2p(s, alpha))) return rc;
	if (T) cudaEventRecord(s->ev[FS3_T_TOTAL], s->stream);
	s->steps++;
	return FS3_OK;
}

// One Simulate(dt) of a lockstep group: phase k of every rank is queued before phase k+1 of any.
int fs3_group_step(fs3_sim** h, int world, float dt) {
	if (!h || world < 2 || !h[0] || world != h[0]->world || !h[0]->lockstep) return FS3_ERR_INVALID;
	cudaSetDevice(h[0]->device);
	dt = effective_dt(h[0], dt);
	const int nph = fs3i_num_step_phases(h[0]);
	for (int ph = 0; ph < nph; ph++)
		for (int r = 0; r < world; r++) { int rc = fs3i_step_phase(h[r], ph, dt); if (rc) return rc; }
	for (int r = 0; r < world; r++) h[r]->steps++;
	return FS3_OK;
}

int fs3_stage_times(fs3_sim* s, float* ms, int cap) {
	ENTER(s);
	if (!ms || cap < FS3_T_COUNT) return FS3_ERR_INVALID;
	if (!s->prm.record_stage_times) return fail(s, FS3_ERR_STATE, "record_stage_times is off");
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	for (int k = 0; k < FS3_T_TOTAL; k++) FS3_CUDA(cudaEventElapsedTime(&ms[k], s->ev[k], s->ev[k + 1]));
	FS3_CUDA(cudaEventElapsedTime(&ms[FS3_T_TOTAL], s->ev[0], s->ev[FS3_T_TOTAL]));
	// the sweeps ran on stream2 concurrently with EXTRAPOLATE: their time is their own span (the stage times then add up
	// to more than TOTAL), and the seed ends where P2G begins
	if (s->ls_overlapped) {
		FS3_CUDA(cudaEventElapsedTime(&ms[FS3_T_LEVELSET_SWEEP], s->ev[FS3_T_LEVELSET_SWEEP], s->ev_ls_end));
		if (!s->ls_seed_forked) FS3_CUDA(cudaEventElapsedTime(&ms[FS3_T_LEVELSET_SEED], s->ev[FS3_T_LEVELSET_SEED], s->ev[FS3_T_P2G]));
		else FS3_CUDA(cudaEventElapsedTime(&ms[FS3_T_BIN], s->ev[FS3_T_BIN], s->ev[FS3_T_P2G]));
		FS3_CUDA(cudaEventElapsedTime(&ms[FS3_T_P2G], s->ev[FS3_T_P2G], s->ev[FS3_T_EXTRAPOLATE]));
	}
	return FS3_OK;
}

// ---- state access -------------------------------------------------------------------------------------
int fs3_get_particles(fs3_sim* s, fs3_particle* host) {
	ENTER(s);
	if (!host) return FS3_ERR_INVALID;
	long long total = s->np;
	if (s->world > 1) { total = 0; for (int r = 0; r < s->world; r++) total += s->np_all[r]; }
	if (total <= 0) return FS3_OK;
	int rc = fs3i_staging(s, (size_t)total * sizeof(fs3_particle));
	if (rc) return rc;
	// sharded: records of particles owned by other ranks read back as NaN (all bits set); the caller merges the ranks
	if (s->world > 1) FS3_CUDA(cudaMemsetAsync(s->stage_ptr, 0xff, (size_t)total * sizeof(fs3_particle), s->stream));
	if ((rc = fs3i_unsort_particles(s, (fs3_particle*)s->stage_ptr))) return rc;
	FS3_CUDA(cudaMemcpyAsync(host, s->stage_ptr, (size_t)total * sizeof(fs3_particle), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_set_particles(fs3_sim* s, const fs3_particle* host, long long count) {
	ENTER(s);
	if (count < 0 || (count > 0 && !host) || count > 0x7fffffffLL) return FS3_ERR_INVALID;
	int rc = s->world > 1 ? FS3_OK : ensure_particle_capacity(s, count > 0 ? count : 1);
	if (rc) return rc;
	if ((rc = fs3i_staging(s, (size_t)(count > 0 ? count : 1) * sizeof(fs3_particle)))) return rc;
	if (count > 0) FS3_CUDA(cudaMemcpyAsync(s->stage_ptr, host, (size_t)count * sizeof(fs3_particle), cudaMemcpyHostToDevice, s->stream));
	// sharded: every rank is given the same full array and keeps the particles of its slab
	rc = s->world > 1 ? fs3i_load_particles_owned(s, (const fs3_particle*)s->stage_ptr, count) : fs3i_load_particles(s, (const fs3_particle*)s->stage_ptr, count);
	if (rc) return rc;
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
// The particles this rank owns, compact, with their creation ids (= index into the reference's m_particles).  The cheap
// per-rank way in and out of a sharded simulation: every rank moves only its own 1/world of the particles.
int fs3_get_local_particles(fs3_sim* s, fs3_particle* host, int* ids, long long cap, long long* count) {
	ENTER(s);
	if (count) *count = s->np;
	if (s->np <= 0) return FS3_OK;
	if (!host || !ids || cap < s->np) return FS3_ERR_INVALID;
	const size_t n = (size_t)s->np;
	int rc = fs3i_staging(s, n * (sizeof(fs3_particle) + 4));
	if (rc) return rc;
	fs3_particle* dp = (fs3_particle*)s->stage_ptr;
	int* di = (int*)((char*)s->stage_ptr + n * sizeof(fs3_particle));
	if ((rc = fs3i_pack_local(s, dp, di))) return rc;
	FS3_CUDA(cudaMemcpyAsync(host, dp, n * sizeof(fs3_particle), cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaMemcpyAsync(ids, di, n * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
// Replaces this rank's particles.  They need not lie in the rank's slab: the next step's migration delivers strays.
int fs3_set_local_particles(fs3_sim* s, const fs3_particle* host, const int* ids, long long count) {
	ENTER(s);
	if (count < 0 || (count > 0 && (!host || !ids))) return FS3_ERR_INVALID;
	if (s->world == 1) return fail(s, FS3_ERR_STATE, "fs3_set_local_particles is for sharded handles (use fs3_set_particles)");
	if (count > s->cap) return fail(s, FS3_ERR_STATE, "more particles than the slab's capacity");
	const size_t n = (size_t)(count > 0 ? count : 1);
	int rc = fs3i_staging(s, n * (sizeof(fs3_particle) + 4));
	if (rc) return rc;
	fs3_particle* dp = (fs3_particle*)s->stage_ptr;
	int* di = (int*)((char*)s->stage_ptr + n * sizeof(fs3_particle));
	if (count > 0) {
		FS3_CUDA(cudaMemcpyAsync(dp, host, (size_t)count * sizeof(fs3_particle), cudaMemcpyHostToDevice, s->stream));
		FS3_CUDA(cudaMemcpyAsync(di, ids, (size_t)count * 4, cudaMemcpyHostToDevice, s->stream));
	}
	if ((rc = fs3i_load_local(s, dp, di, count))) return rc;
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_velocity(fs3_sim* s, float* U, float* V, float* W) {
	ENTER(s);
	{ int rc = fs3i_field_wait(s); if (rc) return rc; }
	if (U) FS3_CUDA(cudaMemcpyAsync(U, s->U, s->nU * 4, cudaMemcpyDeviceToHost, s->stream));
	if (V) FS3_CUDA(cudaMemcpyAsync(V, s->V, s->nV * 4, cudaMemcpyDeviceToHost, s->stream));
	if (W) FS3_CUDA(cudaMemcpyAsync(W, s->W, s->nW * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_set_velocity(fs3_sim* s, const float* U, const float* V, const float* W) {
	ENTER(s);
	if (U) FS3_CUDA(cudaMemcpyAsync(s->U, U, s->nU * 4, cudaMemcpyHostToDevice, s->stream));
	if (V) FS3_CUDA(cudaMemcpyAsync(s->V, V, s->nV * 4, cudaMemcpyHostToDevice, s->stream));
	if (W) FS3_CUDA(cudaMemcpyAsync(s->W, W, s->nW * 4, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_phi(fs3_sim* s, float* phi) {
	ENTER(s);
	if (!phi) return FS3_ERR_INVALID;
	{ int rc = fs3i_field_wait(s); if (rc) return rc; }
	FS3_CUDA(cudaMemcpyAsync(phi, s->phi, s->nC * 4, cudaMemcpyDeviceToHost, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_set_phi(fs3_sim* s, const float* phi) {
	ENTER(s);
	if (!phi) return FS3_ERR_INVALID;
	FS3_CUDA(cudaMemcpyAsync(s->phi, phi, s->nC * 4, cudaMemcpyHostToDevice, s->stream));
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	return FS3_OK;
}
int fs3_get_params(const fs3_sim* s, fs3_params* p) {
	if (!s || !p) return FS3_ERR_INVALID;
	*p = s->prm;
	return FS3_OK;
}
int fs3_set_params(fs3_sim* s, const fs3_params* p) {
	ENTER(s);
	if (!p) return FS3_ERR_INVALID;
	if (p->sor_iterations < 0 || p->sor_iterations > s->res_cap) return fail(s, FS3_ERR_INVALID, "sor_iterations out of range [0,4096]");
	if (p->omega >= 2.0) return fail(s, FS3_ERR_INVALID, "omega must be < 2");
	if (!(p->rho > 0.0f)) return fail(s, FS3_ERR_INVALID, "rho must be positive");
	if (p->levelset_mode != 0 && p->levelset_mode != 1) return fail(s, FS3_ERR_INVALID, "levelset_mode must be 0 or 1");
	if (p->extrapolation_mode != 0 && p->extrapolation_mode != 1) return fail(s, FS3_ERR_INVALID, "extrapolation_mode must be 0 or 1");
	if (p->levelset_mode == 1 && s->world > 1) return fail(s, FS3_ERR_STATE, "the line-sweep level set (levelset_mode 1) is not available on a sharded handle");
	// the slab-window extrapolation of a sharded handle is exact for the nodes the step consumes as long as fluid cells lie within
	// one cell of a particle (p2g.cu, FS3_EX_DEPTH)
	if (s->world > 1 && !(p->particle_radius <= 1.0f)) return fail(s, FS3_ERR_STATE, "particle_radius > 1 is not available on a sharded handle (extrapolation window)");
	FS3_CUDA(cudaStreamSynchronize(s->stream));
	s->prm = *p;
	return fs3i_alloc_pressure(s);
}

}  // extern "C"
