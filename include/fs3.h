/* fs3.h -- C ABI of libfluidsim3_b200.so: the B200-native (sm_100a CUDA) replacement for the
 * reference CPU solver `FluidSim3` (reference: FluidSimulationDemo/Simulation3D.h:18-24,160-162).
 *
 * The reference has no FFI layer; its boundary is the C++ class surface used by its only caller
 * (FluidSimDemoOld3D.cpp:65-67,115,136,234,263-269,437-439).  Each entry point below names the
 * reference member it replaces.  include/FluidSim3Compat.hpp wraps this ABI in a class with the
 * reference's exact names (FluidSim3(x,y,z,cpm), ResetSimulation(), Simulate(dt), m_particles).
 *
 * Conventions: plain pointers and sizes, caller-owned HOST buffers unless a name says `_device`;
 * every call returns FS3_OK (0) or a negative error code and records a message retrievable with
 * fs3_last_error().  Calls on one handle must be serialised by the caller (the reference is
 * single-threaded and not re-entrant, Simulation3D.cpp:107).  There is no CPU fallback: every
 * entry point fails with FS3_ERR_CUDA when no sm_100-class device is usable.
 *
 * Grid layouts are the reference's (Simulation3D.h:41-52), x fastest:
 *   U[(nx+1)*ny*nz]  U(i,j,k) = U[i + (nx+1)*(j + ny*k)]      (x-faces)
 *   V[nx*(ny+1)*nz]  V(i,j,k) = V[i + nx*(j + (ny+1)*k)]      (y-faces)
 *   W[nx*ny*(nz+1)]  W(i,j,k) = W[i + nx*(j + ny*k)]          (z-faces)
 *   phi[nx*ny*nz]    in cell units (Simulation3D.h:156-158)
 */
#ifndef FS3_H
#define FS3_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define FS3_OK 0
#define FS3_ERR_INVALID (-1)    /* bad argument / unsupported shape            */
#define FS3_ERR_CUDA (-2)       /* CUDA runtime error (message in last_error)  */
#define FS3_ERR_STATE (-3)      /* call order / missing prerequisite stage     */
#define FS3_ERR_PARTICLE (-4)   /* particle outside the simulation box         */
#define FS3_ERR_IO (-5)         /* state file missing, truncated or corrupt    */

/* Wire layout of one particle == reference `struct Particle3` (ParticleDefs.h:39-47): 24 bytes. */
typedef struct fs3_particle {
	float X, Y, Z;    /* metres, unit box */
	float uX, uY, uZ; /* m/s */
} fs3_particle;

typedef struct fs3_sim fs3_sim; /* opaque handle == one FluidSim3 object */

/* Runtime parameters; defaults are the reference's hard-coded values. */
typedef struct fs3_params {
	int sor_iterations;     /* 100   (Simulation3D.cpp:946) */
	double omega;           /* <=0 -> fitted 2 - 3.16343/nx (Simulation3D.cpp:944) */
	int precision;          /* 0 = fp64 pressure solve like the reference (default), 1 = fp32 fast mode */
	int force_dt_0_01;      /* 1 = reproduce `dt = 0.01f` override (Simulation3D.cpp:104); 0 = use clamped dt */
	int record_residuals;   /* 1 = record ||b - A p|| after every SOR iteration (instrumentation, slower) */
	float rho, gravity_y, nu, particle_radius; /* 1000, -9.81, 8.9e-7, 1.0 (Simulation3D.h:135-145) */
	int record_stage_times; /* 1 = bracket every stage with CUDA events (fs3_stage_times) */
	int levelset_mode;      /* 0 = ComputeLevelSet of the CPU class: seed + 8 Gauss-Seidel sweeps (default, the parity path);
	                           1 = the reference's GPU formulation: closest particle among 27 cells + 24 line sweeps
	                           (FX/gpComputeClosestParticleNeighbors.hlsl, FX/gpClosestParticlesSweep*.hlsl, Simulation.cpp:718-794).
	                           A different level set (no grid-long dependency chain); single-GPU handles only */
	int extrapolation_mode; /* 0 = ExtrapolateValues of the CPU class: every layer, by L1 distance (default, the parity path);
	                           1 = the reference's GPU formulation: ONE layer, the rest of the air 0
	                           (FX/gpExtrapolateParticleVelocities.hlsl:16-42); on a sharded handle it runs replicated on the
	                           gathered P2G grids, like the default */
} fs3_params;

/* Stage taxonomy for fs3_stage_times, named after the reference profiler marks (GPUProfiler.h:16-44). */
enum {
	FS3_T_ADVECT = 0, FS3_T_BIN, FS3_T_LEVELSET_SEED, FS3_T_LEVELSET_SWEEP, FS3_T_P2G, FS3_T_EXTRAPOLATE,
	FS3_T_PROJECT_SETUP, FS3_T_PROJECT_SOR, FS3_T_PROJECT_TOVEL, FS3_T_FLIP_APPLY, FS3_T_TOTAL, FS3_T_COUNT
};

/* ---- lifecycle -------------------------------------------------------------------------- */
/* FluidSim3::FluidSim3(x,y,z,CellsPerMeter) incl. the ResetSimulation() it runs (Simulation3D.cpp:19-31).
 * Only cubes with cells_per_meter == nx are accepted: the reference mixes mX/mY/mZ and
 * CellsPerMeter (Simulation3D.cpp:283-285 vs :442-444) and indexes uAmts with mZ for mY (:476),
 * so its results are only defined there.  `device` = CUDA ordinal. */
int fs3_create(int nx, int ny, int nz, float cells_per_meter, int device, fs3_sim** out);
/* FluidSim3::~FluidSim3 (Simulation3D.cpp:33-39) */
int fs3_destroy(fs3_sim* s);
/* FluidSim3::ResetSimulation (Simulation3D.cpp:41-99); bit-identical to a libstdc++ build of the
 * reference (minstd_rand0 seed 0, closed-form skip-ahead on the device). */
int fs3_reset(fs3_sim* s);
/* FluidSim3::Simulate(dt) (Simulation3D.cpp:101-188).  Queued on the handle's stream; results are complete after fs3_sync or
 * any getter.  The call itself is NOT free of host waits: the extrapolation reads one bucket size back, a sharded handle reads
 * the world x world migration counts back, and at grid sizes whose pressure unknowns are held in the persisting L2 carve-out
 * (about 184^3 to 320^3 on one GPU) the solve is bracketed by two stream synchronisations and cudaDeviceSetLimit calls (the
 * previous device-wide limit is restored before the call returns).  Do not call it from inside a stream capture. */
int fs3_step(fs3_sim* s, float dt);
/* Blocks until all queued work on the handle is complete and reports deferred device errors. */
int fs3_sync(fs3_sim* s);
const char* fs3_last_error(const fs3_sim* s); /* s may be NULL: error of the last failed fs3_create */

/* ---- z-slab sharding over the GPUs of one box (north_star: "large grids are split into z-slabs") --------------
 * One rank per GPU, one process per rank.  Rank r owns the cell planes z in [r*nz/world, (r+1)*nz/world), the face nodes
 * on them and the particles whose cell lies there.  The reference has no counterpart (single-threaded CPU class); the
 * surface stays FluidSim3's: every rank constructs, steps and reads back with the same calls, which become COLLECTIVE
 * (every rank must make them in the same order).  Ranks talk through CUDA-IPC peer memory over NVLink; the caller only
 * has to carry FS3_SHARD_HANDLE_BYTES per rank between the processes once (MPI_Allgather, torch.distributed, a file):
 *     fs3_create_sharded(...); fs3_shard_get_handle(h, mine); <all-gather the blobs in rank order>; fs3_shard_connect(h, all);
 * fs3_shard_connect ends with the collective ResetSimulation, as the reference constructor does (Simulation3D.cpp:30).
 * Afterwards fs3_step / fs3_reset / fs3_set_particles (same full array on every rank) / fs3_set_velocity / fs3_set_phi work
 * as before; fs3_get_velocity and fs3_get_phi return the whole grids on every rank; fs3_get_particles fills the records
 * of the particles this rank owns and leaves NaN (all bits set) in the others, so the ranks' arrays merge trivially.
 * Every rank needs its own GPU (ranks wait on each other inside kernels); fs3_shard_connect refuses a shared one. */
#define FS3_SHARD_HANDLE_BYTES 128
int fs3_create_sharded(int nx, int ny, int nz, float cells_per_meter, int device, int rank, int world, fs3_sim** out);
int fs3_shard_get_handle(fs3_sim* s, void* blob);
int fs3_shard_connect(fs3_sim* s, const void* blobs_in_rank_order);
int fs3_shard_plan(int nz, int world, int rank, int* z0, int* z1);   /* pure host arithmetic: the slab of a rank */
int fs3_shard_info(const fs3_sim* s, int* rank, int* world, int* z0, int* z1);
long long fs3_num_local_particles(const fs3_sim* s);                   /* particles this rank owns right now */
/* This rank's own particles, compact, with their creation ids (index into the reference's m_particles): the per-rank way in
 * and out, each rank moving 1/world of the data.  Uploaded particles may lie outside the slab; the next step migrates them. */
int fs3_get_local_particles(fs3_sim* s, fs3_particle* host, int* ids, long long cap, long long* count);
int fs3_set_local_particles(fs3_sim* s, const fs3_particle* host, const int* ids, long long count);
/* All ranks inside ONE process on ONE device, sharing a stream and stepped phase by phase ("lockstep"): the same slab
 * kernels, halo stores, migration and ghost-layer code as the multi-process path, but every cross-rank wait is already
 * satisfied by stream order, so it runs (and is parity-tested) on a single GPU.  out/handles = array of `world` handles. */
int fs3_group_create(int nx, int ny, int nz, float cells_per_meter, int device, int world, fs3_sim** out);
int fs3_group_reset(fs3_sim** handles, int world);
int fs3_group_step(fs3_sim** handles, int world, float dt);

/* ---- state access (reference: public m_particles; private m_MU/m_MV/m_MW/m_Phi) ------------ */
long long fs3_num_particles(const fs3_sim* s);                      /* m_particles.size() */
int fs3_get_particles(fs3_sim* s, fs3_particle* host);              /* m_particles, creation order */
int fs3_set_particles(fs3_sim* s, const fs3_particle* host, long long count);
int fs3_get_velocity(fs3_sim* s, float* U, float* V, float* W);     /* any may be NULL */
int fs3_set_velocity(fs3_sim* s, const float* U, const float* V, const float* W);
int fs3_get_phi(fs3_sim* s, float* phi);
int fs3_set_phi(fs3_sim* s, const float* phi);
int fs3_default_params(fs3_params* p);              /* the reference's hard-coded values; host only */
int fs3_get_params(const fs3_sim* s, fs3_params* p);
int fs3_set_params(fs3_sim* s, const fs3_params* p);

/* ---- single stages of Simulate, in call order (parity hooks) ------------------------------ */
int fs3_stage_advect(fs3_sim* s, float dt);       /* Advect, Simulation3D.cpp:190-232 */
int fs3_stage_bin(fs3_sim* s);                    /* counting sort by cell (cell = roundf, :286-288) */
int fs3_stage_levelset(fs3_sim* s);               /* ComputeLevelSet, :255-420 (runs fs3_stage_bin first) */
int fs3_stage_p2g(fs3_sim* s);                    /* TransferParticlesToGrid incl. ExtrapolateValues, :422-778 */
int fs3_stage_bodyforces(fs3_sim* s, float dt);   /* AddBodyForces, :780-788 */
int fs3_stage_project(fs3_sim* s, float dt);      /* Project, :790-1093 (no gravity) */
/* old' = new - (1-alpha) old; u = (1-alpha) u + interp(old')  (:148-180); old grids are host arrays */
int fs3_stage_g2p(fs3_sim* s, float dt, const float* oldU, const float* oldV, const float* oldW);
/* ExtrapolateValues on a caller-supplied grid (:614-778); `valid` is one byte per node */
int fs3_extrapolate(fs3_sim* s, float* grid, const unsigned char* valid, int xs, int ys, int zs);

/* ---- quantities the reference keeps in locals (parity / measurement hooks) ---------------- */
int fs3_get_cell_indices(fs3_sim* s, int* cell);   /* per particle in creation order; valid after bin/levelset */
int fs3_get_cell_counts(fs3_sim* s, int* counts);  /* per cell */
int fs3_get_closest(fs3_sim* s, int* closest);     /* closestParticles[] after the 8 sweeps (creation-order ids) */
int fs3_get_pressure(fs3_sim* s, double* p);       /* p[] of the last Project, natural cell order */
int fs3_get_rhs_diag(fs3_sim* s, double* b, double* diag);
int fs3_get_sor_residuals(fs3_sim* s, double* l2, double* linf, int cap, int* count);
int fs3_stage_times(fs3_sim* s, float* ms, int cap); /* FS3_T_* order, last step, milliseconds */
/* FluidSim3::PrintDivergence (Simulation3D.cpp:1095-1138) as a validation metric: over the cells with phi < 0, the largest
 * (signed) value of (V(y+1)-V(y)) + (U(x+1)-U(x)) + (W(z+1)-W(z)) evaluated in fp32 in that order, and the square root of
 * the sum of its squares accumulated in double.  Sharded handle: the rank's own cell planes only (combine: max of the
 * maxima, square root of the sum of the squared l2 values). */
int fs3_get_divergence(fs3_sim* s, double* max_divergence, double* l2);
/* The same metric WITHOUT any allocation and without a handle.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Inputs are DEVICE pointers to U, V, W, phi in the layouts above (on `device`), the output is a DEVICE pointer to
 * two doubles: out[0] = maximum divergence, out[1] = L2 norm.  Two kernels are queued on `stream` (a cudaStream_t, NULL = the
 * default stream) and the call returns without synchronising; it calls no cudaMalloc / cudaFree or any allocating routine.
 * The block partial sums live in a fixed 1 MiB __device__ array inside the library (held by every GPU that loads the library
 * until the process exits), which limits the grid to FS3_NO_ALLOC_MAX_CELLS cells: larger grids are rejected with
 * FS3_ERR_INVALID (message through fs3_last_error(NULL)).  All calls share that array: calls that overlap on the GPU (different
 * streams or host threads) race -- serialise them on one stream.  The first call on a device looks the array's device address
 * up (cudaGetSymbolAddress); make that call once before starting a graph capture. */
#define FS3_NO_ALLOC_MAX_CELLS (256 * 256 * 256)
int fs3_divergence_no_alloc(int device, void* stream, int nx, int ny, int nz, const float* dev_U, const float* dev_V, const float* dev_W,
                            const float* dev_phi, double* dev_out);
/* Order-independent 64-bit fingerprints of the state, computed on the device (measurement hook: bench.py prints them as
 * `check`, and a sharded run must reproduce the single-GPU values exactly).  Every value is a sum modulo 2^64 of a mixed
 * per-element word, so the ranks' values of a sharded simulation ADD UP to the single-GPU value:
 *   particle_hash: over the rank's particles, mix(creation id, cell index (int)roundf(n*X) (Simulation3D.cpp:286-288),
 *                  bit patterns of X,Y,Z,uX,uY,uZ);  cell_hash: mix(creation id, cell index) only
 *   grid_hash[0..3]: U, V, W, phi over the rank's own planes, mix(element index, bit pattern);  grid_sum: plain sums in double */
typedef struct fs3_check {
	unsigned long long particle_hash, cell_hash;
	long long particles;
	unsigned long long grid_hash[4];
	double grid_sum[4];
} fs3_check;
int fs3_checksum(fs3_sim* s, fs3_check* out);

/* ---- measurement: projection microbench (BASELINE config 3) ------------------------------- */
/* Runs `iters` red-black SOR iterations on the current b/diag/mask (from the last project setup)
 * starting from p = 0 and returns the device time of the sweep loop. */
int fs3_sor_bench(fs3_sim* s, int iters, float* ms_out);
/* Makes every cell fluid (phi = -1, the reference's own debug switch Simulation3D.cpp:851-858). */
int fs3_set_phi_all_fluid(fs3_sim* s);
void* fs3_stream(fs3_sim* s);              /* the cudaStream_t all work is queued on */
long long fs3_kernel_launches(const fs3_sim* s); /* number of kernels launched so far by this handle */

/* ---- phi / binned-particle consumer interface ------------------------------------------
 * What the reference's GPU class hands to its renderer: the public members m_gpPhiSRV, m_gpCountsSRV and
 * m_gpBinnedParticles (Simulation.h:118-140), the phi blur it runs after every step (FX/gpBlur.hlsl:6-17, dispatched
 * at Simulation.cpp:561-565) and the renderer's distance lookup map() (FX/Render.fx:77-115) -- as DEVICE buffers and
 * kernels a downstream CUDA ray-marcher can consume without a graphics API.  Device pointers stay valid until
 * fs3_destroy (the binned buffers: until the next fs3_export_binned); their contents are ordered on fs3_stream().
 * None of these calls changes the solver's state: the blur writes a second buffer, never phi itself. */
int fs3_phi_device(fs3_sim* s, const float** dev_phi);              /* m_gpPhiSRV: phi[nx*ny*nz], cell units */
/* gpBlur: 7-point mean (c + x+ + x- + y+ + y- + z+ + z-) / 7, out-of-range taps read 0 (D3D11 texture semantics).
 * The shader updates phi in place while neighbours read it (racy); this is the deterministic reading: every tap
 * sees the un-blurred phi.  Asynchronous. */
int fs3_blur_phi(fs3_sim* s);
int fs3_phi_blurred_device(fs3_sim* s, const float** dev_phi);     /* result of the last fs3_blur_phi */
int fs3_get_phi_blurred(fs3_sim* s, float* host);
/* m_gpBinnedParticles + prefix-summed m_gpCounts (gpBinParticles.hlsl:13-31, Simulation.cpp:660-686): the particles
 * as 24-byte records grouped by cell (cell = x + nx*(y + ny*z), cell of a particle = roundf(n*X), Simulation3D.cpp:286-288)
 * and offsets[nx*ny*nz + 1] with the records of cell c at [offsets[c], offsets[c+1]).  Order inside a cell: by octant,
 * then creation id (the shader's InterlockedAdd order is unspecified).  Single-GPU handles only. */
int fs3_export_binned(fs3_sim* s, const fs3_particle** dev_binned, const unsigned** dev_offsets);
int fs3_get_binned(fs3_sim* s, fs3_particle* host_binned, unsigned* host_offsets);   /* host copies; either may be NULL */
/* Render.fx map(p): distance to the surface in metres at `count` points (x,y,z triples, unit box) -- quintic-smoothed
 * fractions on x and z, explicit fp32 trilinear filter with clamp addressing in place of the hardware sampler, result
 * scaled by 1/nx.  blurred != 0 samples the fs3_blur_phi buffer.  The _device form takes device pointers and is
 * asynchronous on fs3_stream(). */
int fs3_sample_phi(fs3_sim* s, const float* xyz, long long count, int blurred, float* out);
int fs3_sample_phi_device(fs3_sim* s, const float* dev_xyz, long long count, int blurred, float* dev_out);

/* ---- on-disk state --------------------------------------------------------------------
 * One binary file with everything Simulate() reads at the start of a step (particles in creation order, U, V, W) plus
 * phi, the parameters and the step counter: save, destroy, create, load, and the next fs3_step continues
 * bit-identically.  Replaces the reference's lossy tab-separated text dumps (debugroutines.h:24-178).  Layout:
 * 256-byte header | particles (24 B each) | [creation ids, 4 B each] | U | V | W | phi, little-endian; the header
 * carries a 64-bit checksum of the payload.  Sharded handles: every rank saves / loads its OWN file (the rank's particles
 * with their creation ids + the whole grids) and must be restored with the same rank / world. */
typedef struct fs3_state_info {
	int version;                 /* 1 */
	int nx, ny, nz;
	float cells_per_meter;
	long long num_particles;     /* particle records in this file */
	long long total_particles;   /* particles of the whole simulation (== num_particles unless sharded) */
	long long steps;             /* Simulate() calls since the last ResetSimulation */
	int rank, world;             /* 0, 1 unless the file is one rank of a sharded run */
	int has_ids;                 /* 1 = an id section follows the particles */
	fs3_params params;           /* record_* switches are not stored */
	long long rank_particles[16]; /* live particle count of every rank at save time */
	unsigned long long payload_bytes, checksum;
} fs3_state_info;
int fs3_save_state(fs3_sim* s, const char* path);
int fs3_load_state(fs3_sim* s, const char* path);
long long fs3_step_count(const fs3_sim* s);
/* Host-only file access (no GPU, no handle; errors through fs3_last_error(NULL)): header, whole-file write, whole-file
 * read into caller-owned buffers sized from the header (a NULL buffer skips its section and the checksum test). */
int fs3_state_file_info(const char* path, fs3_state_info* info);
int fs3_state_file_write(const char* path, const fs3_state_info* info, const fs3_particle* particles, const int* ids,
                         const float* U, const float* V, const float* W, const float* phi);
int fs3_state_file_read(const char* path, fs3_state_info* info, fs3_particle* particles, int* ids, float* U, float* V, float* W, float* phi);
unsigned long long fs3_state_checksum(const void* data, size_t bytes, unsigned long long seed);   /* seed 0 starts a chain */

#ifdef __cplusplus
}
#endif
#endif
