// TEST INFRASTRUCTURE ONLY (oracle) -- never linked into the product library.
//
// C-ABI wrapper around the UNMODIFIED reference class FluidSim3
// (/root/reference/FluidSimulationDemo/Simulation3D.{h,cpp}), compiled where the
// sources lie (see oracle/Makefile) into oracle/_ref/libfs3ref.so.  Private state
// and the private per-stage methods are reached with `#define private public`
// around the include -- no reference file is edited or copied.
//
// Used by tests/ to pin the C restatement (oracle/fs3_oracle.c) and to generate
// tests/golden/*; used by bench.py only for the cpu_baseline / --impl reference leg.
#include <vector>
#include <random>
#define private public
#include "Simulation3D.h"
#undef private
#include <cstdint>
#include <cstring>

extern "C" {

void* ref_create(int nx, int ny, int nz, float cpm) { return new FluidSim3(nx, ny, nz, cpm); }
void ref_destroy(void* h) { delete static_cast<FluidSim3*>(h); }
void ref_reset(void* h) { static_cast<FluidSim3*>(h)->ResetSimulation(); }
void ref_step(void* h, float dt) { static_cast<FluidSim3*>(h)->Simulate(dt); }
long ref_num_particles(void* h) { return (long)static_cast<FluidSim3*>(h)->m_particles.size(); }

void ref_get_particles(void* h, float* out6) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	std::memcpy(out6, s->m_particles.data(), s->m_particles.size() * sizeof(Particle3));
}
void ref_set_particles(void* h, const float* in6, long count) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	s->m_particles.resize((size_t)count);
	std::memcpy(s->m_particles.data(), in6, (size_t)count * sizeof(Particle3));
}
static void sizes(FluidSim3* s, size_t& nu, size_t& nv, size_t& nw, size_t& nc) {
	nu = (size_t)(s->mX + 1) * s->mY * s->mZ;
	nv = (size_t)s->mX * (s->mY + 1) * s->mZ;
	nw = (size_t)s->mX * s->mY * (s->mZ + 1);
	nc = (size_t)s->mX * s->mY * s->mZ;
}
void ref_get_velocity(void* h, float* u, float* v, float* w) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	size_t nu, nv, nw, nc; sizes(s, nu, nv, nw, nc);
	if (u) std::memcpy(u, s->m_MU, nu * 4);
	if (v) std::memcpy(v, s->m_MV, nv * 4);
	if (w) std::memcpy(w, s->m_MW, nw * 4);
}
void ref_set_velocity(void* h, const float* u, const float* v, const float* w) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	size_t nu, nv, nw, nc; sizes(s, nu, nv, nw, nc);
	if (u) std::memcpy(s->m_MU, u, nu * 4);
	if (v) std::memcpy(s->m_MV, v, nv * 4);
	if (w) std::memcpy(s->m_MW, w, nw * 4);
}
void ref_get_phi(void* h, float* phi) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	size_t nu, nv, nw, nc; sizes(s, nu, nv, nw, nc);
	std::memcpy(phi, s->m_Phi, nc * 4);
}
void ref_set_phi(void* h, const float* phi) {
	FluidSim3* s = static_cast<FluidSim3*>(h);
	size_t nu, nv, nw, nc; sizes(s, nu, nv, nw, nc);
	std::memcpy(s->m_Phi, phi, nc * 4);
}

// Individual stages of FluidSim3::Simulate (Simulation3D.cpp:110-139), callable one
// at a time so that the CUDA path can be compared stage by stage from identical state.
void ref_stage_advect(void* h, float dt) { FluidSim3* s = static_cast<FluidSim3*>(h); s->Advect(s->m_particles, dt); }
void ref_stage_levelset(void* h) { FluidSim3* s = static_cast<FluidSim3*>(h); s->ComputeLevelSet(s->m_particles); }
void ref_stage_p2g(void* h) { FluidSim3* s = static_cast<FluidSim3*>(h); s->TransferParticlesToGrid(s->m_particles); }
void ref_stage_bodyforces(void* h, float dt) { static_cast<FluidSim3*>(h)->AddBodyForces(dt); }
void ref_stage_project(void* h, float dt) { static_cast<FluidSim3*>(h)->Project(dt); }
void ref_stage_extrapolate(void* h, float* src, const unsigned char* valid, int xs, int ys, int zs) {
	// bool is 1 byte with values 0/1 on this ABI.
	static_cast<FluidSim3*>(h)->ExtrapolateValues(src, (bool*)valid, xs, ys, zs);
}
void ref_interpolate(void* h, float i, float j, float k, float* out3) {
	XMFLOAT3 r = static_cast<FluidSim3*>(h)->InterpolateMACCell(i, j, k);
	out3[0] = r.x; out3[1] = r.y; out3[2] = r.z;
}

}  // extern "C"
