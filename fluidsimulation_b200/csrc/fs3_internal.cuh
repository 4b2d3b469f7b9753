// Internal state and helpers shared by the sm_100a kernels of libfluidsim3_b200.so.
// Reference citations "S3D.cpp:a-b" / "S3D.h:a-b" are FluidSimulationDemo/Simulation3D.{cpp,h}.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include "../../include/fs3.h"

#define FS3_INF_CD 1000000000  // ExtrapolateValues' "inf" (S3D.cpp:632)
#define FS3_MAX_WORLD 16
// words of a rank's flag block `bar`: [0,16) barrier epochs by rank; [16,18) SOR epochs of the lower / upper neighbour;
// [32,48) "velocity halo pushed" epochs by rank; [48,64) "whole slab pushed" epochs by rank
#define FS3_BAR_WORDS (4 * FS3_MAX_WORLD)
#define FS3_BAR_XNEAR (2 * FS3_MAX_WORLD)
#define FS3_BAR_XALL (3 * FS3_MAX_WORLD)

// ---- red-black ("colour-split") storage of the pressure system ---------------------------------
// Cells with (x+y+z)&1 == c live in colour array c at  xh + hp*(y + ny*(z+1)),  xh = x>>1.
// hp (half-row pitch) is a multiple of 4 so that every row starts 16-byte aligned for fp32 and fp64;
// one ghost plane below z=0 and one above z=nz-1 make the 7-point stencil branch-free (and are the
// z-slab halos in the multi-GPU layout).
struct RBLayout {
	int nx, ny, nz;
	int hx;            // colour cells per row = (nx+1)/2
	int hp;            // pitch
	size_t plane;      // hp*ny
	size_t total;      // plane*(nz+2)
	__host__ __device__ size_t at(int xh, int y, int z) const { return (size_t)xh + (size_t)hp * ((size_t)y + (size_t)ny * (size_t)(z + 1)); }
};

struct fs3_sim {
	int device;
	int nx, ny, nz;
	float cpm;
	fs3_params prm;
	cudaStream_t stream;
	std::string err;
	long long launches;
	long long steps;         // Simulate() calls since the last ResetSimulation (saved in state files)

	// MAC grids.  cur = field seen by Advect / returned by get_velocity (the projected field);
	// old = target of P2G, kept as "old grid" for FLIP (S3D.cpp:123-156).
	size_t nU, nV, nW, nC;
	float *U, *V, *W;
	float *oU, *oV, *oW;
	float* phi;

	// particles, SoA of two float4: pos = (X,Y,Z, bits(creation id)), vel = (uX,uY,uZ,0); double-buffered for the sort
	long long np;
	long long cap;
	float4 *pos[2], *vel[2];
	int cur;                 // which buffer holds the live particles
	bool sorted;             // live buffer is half-cell-sorted and cell_start is valid
	bool counted;            // Advect already produced cellid / arrival / cell_count of the live buffer (single GPU: fused)
	size_t nH;               // 8*nC half-cells: the sort key is the particle's octant inside its cell, (2n)^3 grid, x fastest
	uint32_t* cellid;        // per particle (live order), cell index
	uint32_t* cell_count;    // nH+1
	uint32_t* cell_start;    // nH+1 exclusive scan
	uint32_t* cell_fill;     // per particle (live order): arrival order inside its half-cell (nH words >= any particle count)
	int* tmp_id;             // per sorted slot: creation id (ranking key)
	uint32_t* scan_block;    // look-back scan scratch: scan_block_cap 64-bit tile states, then the ticket
	size_t scan_block_cap;

	// level set
	float4* ls_rec;          // per cell: closest particle (n*X, n*Y, n*Z, bits(id)); id<0 = none
	int* closest;            // per cell creation id (exported)
	bool closest_valid;
	// persistent dataflow sweep: tiles in wavefront order, per-tile completion epochs, ticket counter
	uint32_t* ls_order;      // 8*ntiles work items (Tx | Ty<<9 | Tz<<18 | sweep<<27), sorted by DAG level
	uint32_t* ls_done;       // ntiles: epoch of the last sweep this tile completed
	uint32_t* ls_ticket;     // 1 word
	uint32_t ls_epoch;       // host counter: epoch of sweep m of the current call = ls_epoch + m + 1
	int ls_sleep_ns;         // cap of the poll back-off
	int ls_ntiles, ls_blocks, ls_levels;   // ls_levels = depth of the dependency DAG (critical path in tiles)
	long long ls_nitems;     // (sweep, tile) items this handle's sweep kernel runs

	// extrapolation scratch (sized for the largest face grid)
	int* cdv[3];             // L1 distance, one slot per face grid
	int* orderv[3];          // nodes with cd>0 bucketed by distance
	unsigned char* exmask[3]; // per list entry: which of the node's 6 neighbours are strictly closer
	uint32_t* bucket;        // per slot: histogram / starts / cursors (3*bucket_cap+2 words)
	int bucket_cap;
	uint32_t* ex_arrive;     // arrive counter of the layered fill's grid barrier

	// pressure system
	RBLayout rb;
	void *p[2], *b[2], *dg[2]; // colour arrays, double or float according to prm.precision
	int alloc_precision;
	unsigned char* mask[2];     // bit0..5: x-,x+,y-,y+,z-,z+ neighbour is fluid; bit6: cell is fluid
	bool system_valid;
	// active-vector lists of the two colours (rebuilt by every projection setup)
	uint32_t* sor_flags;        // nvec+1
	uint32_t* sor_pos[2];       // nvec+1: exclusive scan of the flags; [nvec] = number of active vectors
	uint32_t* sor_list[2];      // compacted: vector index | row parity << 31
	uint32_t* sor_count[2];     // = sor_pos[c] + nvec (device pointer to the list length)
	bool sor_list_valid;
	int sm_count;
	int sor_per_sm;            // resident CTAs per SM of the sweep kernel (0 = not queried yet)
	long long persist_cap;     // largest persisting-L2 carve-out of the device (-1 = not queried yet)
	double* res;               // device: [2*i] sum r^2, [2*i+1] max |r| (as double bits)
	int res_cap, res_count;

	// ---- z-slab sharding over the GPUs of one box (world == 1: the whole grid) ------------------------------------
	// Every rank keeps FULL-SIZE grid arrays with global indexing and works on the cells / face nodes with z in
	// [z0, z1); particles are owned by the rank whose slab contains their cell.  All device memory of a sharded
	// handle comes from ONE arena, so one CUDA-IPC handle per rank makes every array of every rank addressable
	// from every rank at the same offset (fs3_peer).  Ranks exchange data with plain stores / copies into the
	// peers' arrays over NVLink and order them with epoch flags (shard.cu).
	int rank, world, z0, z1;
	bool lockstep;             // all ranks live in this process, share one stream and are stepped phase by phase:
	                           // waits are then satisfied by stream order and never spin (single-GPU testing)
	bool owns_stream;
	char* arena;
	size_t arena_bytes, arena_used;
	char* peer_base[FS3_MAX_WORLD];   // arena of rank r mapped into this process (own rank: arena)
	int n_ipc_opened;
	size_t poff;               // live particles sit at [poff, poff+np_in) of pos/vel; ghost layers just below / above
	long long gcap;            // capacity of each ghost region (== poff)
	long long np_in;           // length of the live range including dead slots (id < 0) left by emigrants; == np when sorted
	long long np_all[FS3_MAX_WORLD];  // live particle count of every rank, tracked identically on all ranks
	uint32_t* mig_matrix;      // world x (world+1): row r (written by rank r into every copy) = particles r sends to each rank, then the length of r's live range
	uint32_t* mig_ctl;         // [0,world) counts, [world,2world) cursors, [2world] number of emigrants
	uint32_t* leaver_idx;      // indices (relative to poff) of this step's emigrants
	float4 *mig_pos, *mig_vel; // emigrants grouped by destination
	long long mig_cap;
	uint32_t* h_pin;           // pinned host scratch (count matrix)
	uint32_t* bar;             // [r] barrier epoch written by rank r; [FS3_MAX_WORLD+0/1] SOR epoch of the lower/upper neighbour
	uint32_t bar_epoch, sor_epoch;
	uint32_t xfer_epoch;       // steps whose velocity halo this rank has pushed (identical on all ranks)
	uint32_t gather_epoch;     // whole-slab gathers made for the whole-grid getters (identical on all ranks: the getters are collective)
	bool field_dirty;          // a step has run since the last gather
	uint32_t* sor_done;        // block counter of the fused end-of-sweep signal
	int sor_halves;            // half sweeps of the current fs3i_sor_loop call
	bool ls_seed_forked;       // the seed of the current step ran on stream2 in front of the sweeps
	float sor_persist_ratio;   // > 0 while the persisting L2 carve-out is up: hit ratio of the half sweeps' access window
	uint32_t sor_e0;           // SOR epoch of "system set up" of the current projection

	// consumer interface (consumer.cu): allocated on first use, never touched by Simulate
	float* phi_blur;           // 7-point mean of phi (gpBlur.hlsl) as of the last fs3_blur_phi
	bool phi_blur_valid;
	uint32_t* bin_offsets;     // nC+1: cell-major exclusive offsets into `binned`
	fs3_particle* binned;      // particles as 24-byte records, cell by cell
	size_t phi_blur_bytes, bin_offsets_bytes, binned_bytes;

	int* dev_flags;            // [0] particle outside, [1] level-set wait timeout, [2] cross-rank wait failed, [3] shard capacity,
	                           // [4] Advect left the velocity halo of a sharded handle (8 words)
	void* stage_ptr;           // device staging for AoS / hook transfers at the boundary
	size_t stage_bytes;
	cudaEvent_t ev[FS3_T_COUNT + 2];
	// the 8 level-set sweeps are latency-bound and independent of P2G + extrapolation: inside a step they run on a second stream
	cudaStream_t stream2;    // level-set sweeps
	cudaStream_t stream3;    // sharded handles: slab all-gathers (copies + the barrier signal that follows them)
	cudaEvent_t ev_fork, ev_join, ev_ls_end, ev_xfer;
	bool ls_forked;          // sweeps of the current step are in flight on stream2
	bool ls_overlapped;      // the last step's LEVELSET_SWEEP time is (ev[LEVELSET_SWEEP], ev_ls_end) on stream2
	float stage_ms[FS3_T_COUNT];
};

#define FS3_CUDA(call)                                                                           \
	do {                                                                                         \
		cudaError_t e_ = (call);                                                                 \
		if (e_ != cudaSuccess) {                                                                 \
			s->err = std::string(#call) + ": " + cudaGetErrorString(e_);                         \
			return FS3_ERR_CUDA;                                                                 \
		}                                                                                        \
	} while (0)

#define FS3_LAUNCH_CHECK()                                                                       \
	do {                                                                                         \
		s->launches++;                                                                           \
		cudaError_t e_ = cudaGetLastError();                                                     \
		if (e_ != cudaSuccess) {                                                                 \
			s->err = std::string("kernel launch: ") + cudaGetErrorString(e_);                    \
			return FS3_ERR_CUDA;                                                                 \
		}                                                                                        \
	} while (0)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- device helpers common to the particle kernels ---------------------------------------------
struct GridDims {
	int nx, ny, nz;
};

__device__ __forceinline__ float fs3_clamp(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }  // MathHelper.h:49-53
__device__ __forceinline__ float fs3_lerp(float a, float b, float t) { return __fadd_rn(a, __fmul_rn(__fsub_rn(b, a), t)); }  // MathHelper.h:43-47, never fused

// One axis of InterpolateMACCell (S3D.h:61-82): clamp, floor, step back from the last node, fraction.
__device__ __forceinline__ void fs3_split(float c, float hi, int last, int& i0, float& f) {
	float cc = fs3_clamp(c, 0.0f, hi);
	int i = (int)floorf(cc);
	if (i == last) i--;
	i0 = i;
	f = __fsub_rn(cc, (float)i);
}

// Four row pointers (x run of 2 at y, y+1 on planes z, z+1) and eight loads at immediate offsets 0 / 4: the address
// arithmetic is 4 pointer adds per component instead of one per tap (the particle kernels are issue-bound).  Element indices
// fit in 32 bits for every grid the library accepts (the half-cell keys already require (2n)^3 < 2^32).
__device__ __forceinline__ float fs3_tri(const float* __restrict__ g, uint32_t base, uint32_t sy, uint32_t sz, float fx, float fy, float fz) {
	// x lerps first, then y, then z (S3D.h:87-97).  Row starts as 32-bit element indices (one add each), one widening
	// multiply-add per row pointer.
	const uint32_t b10 = base + sy, b01 = base + sz, b11 = b01 + sy;
	const float* __restrict__ r00 = g + base;
	const float* __restrict__ r10 = g + b10;
	const float* __restrict__ r01 = g + b01;
	const float* __restrict__ r11 = g + b11;
	float a00 = fs3_lerp(__ldg(r00), __ldg(r00 + 1), fx);
	float a10 = fs3_lerp(__ldg(r10), __ldg(r10 + 1), fx);
	float a01 = fs3_lerp(__ldg(r01), __ldg(r01 + 1), fx);
	float a11 = fs3_lerp(__ldg(r11), __ldg(r11 + 1), fx);
	return fs3_lerp(fs3_lerp(a00, a10, fy), fs3_lerp(a01, a11, fy), fz);
}

// InterpolateMACCell (S3D.h:55-123) at index-space point (i,j,k).
__device__ __forceinline__ float3 fs3_interp(const float* __restrict__ U, const float* __restrict__ V, const float* __restrict__ W,
                                             GridDims d, float i, float j, float k) {
	int iI, iJ, iK, eI, eJ, eK;
	float fI, fJ, fK, gI, gJ, gK;
	fs3_split(i, (float)d.nx - 1.0f, d.nx - 1, iI, fI);
	fs3_split(j, (float)d.ny - 1.0f, d.ny - 1, iJ, fJ);
	fs3_split(k, (float)d.nz - 1.0f, d.nz - 1, iK, fK);
	fs3_split(__fadd_rn(i, 0.5f), (float)d.nx, d.nx, eI, gI);
	fs3_split(__fadd_rn(j, 0.5f), (float)d.ny, d.ny, eJ, gJ);
	fs3_split(__fadd_rn(k, 0.5f), (float)d.nz, d.nz, eK, gK);
	const uint32_t nx = (uint32_t)d.nx, ny = (uint32_t)d.ny;
	float3 r;
	r.x = fs3_tri(U, (uint32_t)eI + (nx + 1u) * ((uint32_t)iJ + ny * (uint32_t)iK), nx + 1u, (nx + 1u) * ny, gI, fJ, fK);
	r.y = fs3_tri(V, (uint32_t)iI + nx * ((uint32_t)eJ + (ny + 1u) * (uint32_t)iK), nx, nx * (ny + 1u), fI, gJ, fK);
	r.z = fs3_tri(W, (uint32_t)iI + nx * ((uint32_t)iJ + ny * (uint32_t)eK), nx, nx * ny, fI, fJ, gK);
	return r;
}

// cell of a particle: (int)roundf(n*X) per axis (S3D.cpp:283-288)
__device__ __forceinline__ int fs3_cell_axis(float p, float n) { return (int)roundf(__fmul_rn(p, n)); }
// half-cell of a particle on one axis: 2*cell + (n*X >= cell centre).  Half-cell h spans [h/2 - 0.5, h/2) in
// index space, so the support of any trilinear-hat node is exactly 4 consecutive half-cells per axis.
__device__ __forceinline__ int fs3_halfcell_axis(float p, float n, int& cell) {
	float q = __fmul_rn(p, n);
	float r = roundf(q);
	cell = (int)r;
	return 2 * cell + (q >= r ? 1 : 0);
}
// particles of cell (cx,cy,cz) = 4 half-rows of 2 consecutive half-cells: visit(range begin, range end)
#define FS3_FOR_CELL_RANGES(start, d, cx, cy, cz, B, E)                                                          \
	for (int hr_ = 0; hr_ < 4; hr_++)                                                                            \
		for (uint32_t hq_ = 0, B = __ldg((start) + (size_t)(2 * (cx)) + (size_t)(2 * (d).nx) * ((size_t)(2 * (cy) + (hr_ & 1)) + (size_t)(2 * (d).ny) * (2 * (cz) + (hr_ >> 1)))), \
		              E = __ldg((start) + (size_t)(2 * (cx)) + 2 + (size_t)(2 * (d).nx) * ((size_t)(2 * (cy) + (hr_ & 1)) + (size_t)(2 * (d).ny) * (2 * (cz) + (hr_ >> 1)))); hq_ < 1; hq_++)

// ---- cross-rank epoch flags (system scope: the writer is another GPU) ---------------------------------
__device__ __forceinline__ uint32_t fs3_ld_acquire_sys(const uint32_t* p) {
	uint32_t v;
	asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ uint32_t fs3_ld_relaxed_sys(const uint32_t* p) {
	uint32_t v;
	asm volatile("ld.relaxed.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void fs3_st_relaxed_sys(uint32_t* p, uint32_t v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void fs3_st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory"); }
// Waits until *f has reached `need` (epochs wrap: signed difference).  Never hangs: gives up after ~4 s (or at once in
// lockstep mode, where stream order must already have satisfied the wait) and raises flags[2].
__device__ __forceinline__ void fs3_flag_wait(const uint32_t* f, uint32_t need, int* flags, int nospin) {
	if ((int32_t)(fs3_ld_acquire_sys(f) - need) >= 0) return;
	if (nospin) { flags[2] = 1; return; }
	long long t0 = clock64();
	for (unsigned it = 1;; it++) {
		if ((int32_t)(fs3_ld_relaxed_sys(f) - need) >= 0) break;
		__nanosleep(64);
		if ((it & 1023u) == 0) {
			if (*reinterpret_cast<volatile int*>(flags + 2)) return;
			if (clock64() - t0 > 8000000000LL) { flags[2] = 1; return; }
		}
	}
	(void)fs3_ld_acquire_sys(f);
}


// ---- host-side launchers implemented in the .cu files -------------------------------------------
int fs3i_reset(fs3_sim* s);
int fs3i_advect(fs3_sim* s, float dt);
int fs3i_bin(fs3_sim* s);
int fs3i_scan_u32(fs3_sim* s, const uint32_t* in, uint32_t* out, size_t n, uint32_t base0 = 0);   // exclusive scan (+ base0) on the handle's stream
int fs3i_unsort_particles(fs3_sim* s, fs3_particle* dev_out);              // creation order AoS into device buffer
int fs3i_load_particles(fs3_sim* s, const fs3_particle* dev_in, long long count);
int fs3i_cell_indices(fs3_sim* s, int* dev_out);                            // creation order
int fs3i_levelset(fs3_sim* s);
int fs3i_p2g(fs3_sim* s);
int fs3i_extrapolate(fs3_sim* s, float* grid, int xs, int ys, int zs);          // uses slot 0 (cdv[0] initialised)
int fs3i_cd_from_valid(fs3_sim* s, const unsigned char* dev_valid, size_t n);
int fs3i_bodyforces(fs3_sim* s, float dt);
int fs3i_project(fs3_sim* s, float dt, bool from_old, bool with_gravity);
int fs3i_sor_loop(fs3_sim* s, int iters);
int fs3i_clear_pressure(fs3_sim* s);
int fs3i_export_pressure(fs3_sim* s, double* dev_p, double* dev_b, double* dev_dg);
int fs3i_flip_combine(fs3_sim* s, float alpha);                             // old <- cur - (1-alpha) old
int fs3i_g2p(fs3_sim* s, float alpha);
int fs3i_alloc_pressure(fs3_sim* s);
int fs3i_levelset_init(fs3_sim* s);
int fs3i_staging(fs3_sim* s, size_t bytes);
int fs3i_extrap_prepare(fs3_sim* s, int slot, int xs, int ys, int zs, size_t off = 0);
#define FS3_EX_DEPTH 8
#define FS3_SOR_DONE_BYTES 8192   // sor_done: block counter, then (128 bytes apart) arrivals + generation copies of the solve's grid barrier                                                        // sharded handles: planes of P2G halo and layers of extrapolation (p2g.cu)
void fs3i_ex_window(const fs3_sim* s, int ax, int depth, int* w0, int* w1);  // planes [w0, w1) of face grid ax this rank extrapolates
int fs3i_extrap_fill(fs3_sim* s, int slot, float* grid, int xs, int ys, int zs);
int fs3i_alloc(fs3_sim* s, void** p, size_t bytes);                         // arena (sharded) or cudaMalloc
void fs3i_free(fs3_sim* s, void* p);
int fs3i_check_flags(fs3_sim* s);
// particles.cu (sharded pieces)
int fs3i_reset_grids(fs3_sim* s);
int fs3i_reset_particles_slab(fs3_sim* s);
int fs3i_mig_mark(fs3_sim* s);                                              // emigrants of the live range (not fused with Advect)
int fs3i_mig_pack(fs3_sim* s);
int fs3i_load_particles_owned(fs3_sim* s, const fs3_particle* dev_in, long long count);
int fs3i_pack_local(fs3_sim* s, fs3_particle* dev_out, int* dev_ids);
int fs3i_load_local(fs3_sim* s, const fs3_particle* dev_in, const int* dev_ids, long long count);
// project.cu (sharded pieces)
int fs3i_project_setup(fs3_sim* s, float dt, bool from_old, bool with_gravity);
int fs3i_sor_prepare(fs3_sim* s);
int fs3i_sor_half_launch(fs3_sim* s, int k);                                // half sweep k of the current projection
bool fs3i_sor_two_plane(const fs3_sim* s);                                  // sharded solve with two-plane halos (one exchange per iteration)?
int fs3i_sor_push_system(fs3_sim* s);                                       // two-plane halos: b, diag, mask of my boundary reds to the neighbours
int fs3i_sor_persistent(fs3_sim* s, int iters);                             // multi-process sharded handle: the whole solve in one cooperative launch
int fs3i_project_gradient(fs3_sim* s, float dt, bool from_old, bool with_gravity);
// levelset.cu / p2g.cu pieces
int fs3i_levelset_seed(fs3_sim* s, cudaStream_t st = nullptr);
int fs3i_levelset_sweeps(fs3_sim* s, cudaStream_t st = nullptr);                // st: stream to run on (default: the handle's)
int fs3i_levelset_fork(fs3_sim* s, bool after_barrier = false, bool with_seed = false);              // sweeps on stream2, after everything queued so far (and, sharded, after the current barrier epoch)
int fs3i_levelset_join(fs3_sim* s);                                          // main stream waits for them
int fs3i_p2g_gather(fs3_sim* s);
int fs3i_p2g_gather_axis(fs3_sim* s, int ax);                                // one face grid of the slab
int fs3i_cd_from_tags(fs3_sim* s);                                           // sharded: validity of the other slabs' nodes from the gathered grids
void fs3i_p2g_range(const fs3_sim* s, int ax, size_t* q0, size_t* q1);
int fs3i_extrapolate_all(fs3_sim* s);
// shard.cu
int fs3i_shard_alloc(fs3_sim* s);
int fs3i_signal(fs3_sim* s, cudaStream_t st = nullptr);                      // barrier, first half: publish my epoch to all ranks
int fs3i_wait(fs3_sim* s, cudaStream_t st = nullptr);                        // barrier, second half: all ranks reached my epoch
int fs3i_push(fs3_sim* s, void* ptr, size_t byte_off, size_t bytes, cudaStream_t st = nullptr);
int fs3i_flag_signal(fs3_sim* s, unsigned mask, int slot0, uint32_t epoch, cudaStream_t st = nullptr);
int fs3i_flag_wait(fs3_sim* s, unsigned mask, int slot0, uint32_t epoch, cudaStream_t st = nullptr);
int fs3i_field_push(fs3_sim* s, cudaStream_t st);
int fs3i_field_wait(fs3_sim* s);                                            // whole-grid getters: every rank's slab of the projected field has landed
int fs3i_velocity_halo(const fs3_sim* s);   // copy a range of one of my arrays into the same range of every peer
int fs3i_sor_signal(fs3_sim* s);                                            // "system set up": first SOR epoch of this projection to both neighbours
int fs3i_sor_wait(fs3_sim* s, uint32_t epoch);
int fs3i_ghost_pull(fs3_sim* s);
int fs3i_mig_publish(fs3_sim* s);
int fs3i_mig_exchange(fs3_sim* s);
int fs3i_step_phase(fs3_sim* s, int phase, float dt);
int fs3i_reset_phase(fs3_sim* s, int phase);
int fs3i_num_step_phases(const fs3_sim* s);
size_t fs3i_arena_bytes(int n, long long ptot, long long mig_cap, int world);
// consumer.cu
void fs3i_consumer_free(fs3_sim* s);
// levelset_lines.cu
int fs3i_levelset_lines(fs3_sim* s);                                         // fs3_params.levelset_mode == 1

template <typename T> static inline T* fs3_peer(const fs3_sim* s, int r, T* p) {
	return reinterpret_cast<T*>(s->peer_base[r] + (reinterpret_cast<char*>(p) - s->arena));
}
