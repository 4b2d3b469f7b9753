/* TEST INFRASTRUCTURE ONLY -- CPU oracle, never part of the product path.
 *
 * Plain-C restatement of the reference CPU solver FluidSim3
 * (/root/reference/FluidSimulationDemo/Simulation3D.{h,cpp}); every function in
 * fs3_oracle.c cites the reference lines it restates.  Pinned (tests/test_oracle.py)
 * against the reference itself compiled headless into oracle/_ref/libfs3ref.so and
 * against tests/golden/ produced from that build.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  Build with -O2 -ffp-contract=off (the reference is an
 * MSVC /fp:precise build: no FMA contraction).
 */
#ifndef FS3_ORACLE_H
#define FS3_ORACLE_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct fs3o fs3o;

fs3o* fs3o_create(int nx, int ny, int nz, float cells_per_meter); /* ctor + ResetSimulation */
void fs3o_destroy(fs3o* s);
void fs3o_reset(fs3o* s);
void fs3o_step(fs3o* s, float dt); /* FluidSim3::Simulate */
long fs3o_num_particles(const fs3o* s);

void fs3o_get_particles(const fs3o* s, float* out6);          /* AoS X,Y,Z,uX,uY,uZ */
void fs3o_set_particles(fs3o* s, const float* in6, long count);
void fs3o_get_velocity(const fs3o* s, float* u, float* v, float* w);
void fs3o_set_velocity(fs3o* s, const float* u, const float* v, const float* w);
void fs3o_get_phi(const fs3o* s, float* phi);
void fs3o_set_phi(fs3o* s, const float* phi);

/* stage-by-stage entry points (same order as Simulate) */
void fs3o_stage_advect(fs3o* s, float dt);
void fs3o_stage_levelset(fs3o* s);
void fs3o_stage_p2g(fs3o* s);
void fs3o_stage_bodyforces(fs3o* s, float dt);
void fs3o_stage_project(fs3o* s, float dt);
/* old' = new - (1-alpha) old ; u = (1-alpha) u + interp(old')  (Simulation3D.cpp:148-180) */
void fs3o_stage_g2p(fs3o* s, float dt, const float* old_u, const float* old_v, const float* old_w);
void fs3o_extrapolate(float* src, const unsigned char* valid, int xs, int ys, int zs);
void fs3o_interpolate(const fs3o* s, float i, float j, float k, float* out3);

/* parity hooks: quantities the reference computes in locals and throws away */
void fs3o_get_cell_indices(const fs3o* s, int* cell);   /* per particle, -1 if outside */
void fs3o_get_cell_counts(const fs3o* s, int* counts);  /* per cell */
void fs3o_get_closest(const fs3o* s, int* closest);     /* after last levelset stage */
void fs3o_get_pressure(const fs3o* s, double* p);       /* after last project stage */
void fs3o_get_rhs_diag(const fs3o* s, double* b, double* diag); /* diag = 0 in air */
int  fs3o_get_residuals(const fs3o* s, double* l2, double* linf, int cap); /* per iteration */
void fs3o_set_sor_iterations(fs3o* s, int iters);       /* default 100 (Simulation3D.cpp:946) */
void fs3o_set_record_residuals(fs3o* s, int on);
/* fp32 variant of Project (all doubles replaced by float) to quantify fast-mode drift */
void fs3o_stage_project_f32(fs3o* s, float dt);
/* max and L2 of divergence over phi<0 cells, as PrintDivergence (Simulation3D.cpp:1095-1138) */
void fs3o_divergence(const fs3o* s, double* maxdiv, double* l2);

/* ---- consumer interface (fs3_consumer_oracle.c; PARITY UNPINNED: restates D3D11 shaders that cannot run here) ---- */
void fs3o_blur_phi(const float* phi, float* out, int nx, int ny, int nz);                    /* FX/gpBlur.hlsl:6-17 */
void fs3o_sample_phi(const float* phi, int nx, int ny, int nz, const float* xyz, long count, float* out); /* FX/Render.fx:77-115 */
void fs3o_bin_particles(const float* part6, long np, int nx, int ny, int nz, float cpm, float* binned6, unsigned* offsets); /* FX/gpBinParticles.hlsl:13-31 */

/* ---- line-sweep level set (fs3_gpu_semantics_oracle.c; PARITY UNPINNED: restates D3D11 shaders that cannot run here) ---- */
void fs3o_gls_extrapolate(float* grid, const unsigned char* valid, int xs, int ys, int zs);   /* FX/gpExtrapolateParticleVelocities.hlsl:16-42 */
int fs3o_gls_levelset(const float* part6, long np, int nx, int ny, int nz, float cpm, float radius, float* phi, int* closest);

#ifdef __cplusplus
}
#endif
#endif
