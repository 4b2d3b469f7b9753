// FluidSim3Compat.hpp -- header-only C++ host side of libfluidsim3_b200.so.
//
// A class with the reference's exact surface (FluidSimulationDemo/Simulation3D.h:18-24,160-162):
//     FluidSim3(int xSize, int ySize, int zSize, float CellsPerMeter);  ~FluidSim3();
//     void ResetSimulation();   void Simulate(float dt);   std::vector<Particle3> m_particles;
// so that the reference's only caller (FluidSimDemoOld3D.cpp:115,136,234,263-269,437-439) compiles against it
// unchanged: include this header instead of "Simulation3D.h" and link -lfluidsim3_b200.  All work runs on the
// GPU through the C ABI in fs3.h; the header contains no solver arithmetic and there is no CPU fallback
// (construction throws std::runtime_error when no sm_100 device is usable).
//
// m_particles is a public data member in the reference and is read right after Simulate() by the caller
// (FluidSimDemoOld3D.cpp:263-269), so by default it is refreshed (one 24 B/particle device->host copy) at
// the end of every Simulate()/ResetSimulation().  Callers that step many times between reads switch that off
// with SetParticleReadback(false) and call FetchParticles() when they want the vector filled.
#ifndef FS3_FLUIDSIM3_COMPAT_HPP
#define FS3_FLUIDSIM3_COMPAT_HPP
#include <stdexcept>
#include <string>
#include <vector>
#include "fs3.h"

#ifndef NB_FLUIDSIMULATION_PARTICLEDEFS_H   // the reference's own ParticleDefs.h may already be in the TU
// Same 24-byte layout and member names as the reference's Particle3 (ParticleDefs.h:39-60).
struct Particle3 {
	float X, Y, Z;      // metres, unit box
	float uX, uY, uZ;   // m/s
	Particle3() : X(0.f), Y(0.f), Z(0.f), uX(0.f), uY(0.f), uZ(0.f) {}
	Particle3(float px, float py, float pz, float ux, float uy, float uz) : X(px), Y(py), Z(pz), uX(ux), uY(uy), uZ(uz) {}
};
#endif
static_assert(sizeof(Particle3) == sizeof(fs3_particle), "Particle3 must be the 24-byte wire layout");

class FluidSim3 {
public:
	FluidSim3(int xSize, int ySize, int zSize, float CellsPerMeter) : FluidSim3(xSize, ySize, zSize, CellsPerMeter, 0) {}
	// extra argument: CUDA device ordinal
	FluidSim3(int xSize, int ySize, int zSize, float CellsPerMeter, int device) : mX(xSize), mY(ySize), mZ(zSize) {
		int rc = fs3_create(xSize, ySize, zSize, CellsPerMeter, device, &m_h);
		if (rc != FS3_OK) throw std::runtime_error(std::string("fs3_create: ") + fs3_last_error(nullptr));
		FetchParticles();   // the reference constructor ends with ResetSimulation() (Simulation3D.cpp:30)
	}
	~FluidSim3() { fs3_destroy(m_h); }
	FluidSim3(const FluidSim3&) = delete;
	FluidSim3& operator=(const FluidSim3&) = delete;

	void ResetSimulation() {
		Check(fs3_reset(m_h), "fs3_reset");
		m_dirty = false;   // host edits made before the reset are gone, like the vector the reference rebuilds (Simulation3D.cpp:73-98)
		if (m_readback) FetchParticles();
	}
	// The reference clamps dt to [0, 1/15] and then overrides it with 0.01 (Simulation3D.cpp:103-104);
	// the library reproduces both unless fs3_params.force_dt_0_01 is cleared.
	void Simulate(float dt) {
		if (m_dirty) PushParticles();
		Check(fs3_step(m_h, dt), "fs3_step");
		// fs3_sync also reports deferred device errors (a particle outside the box, a dependency wait that timed out, a slab
		// over capacity) that the asynchronous step could not return
		if (m_readback) { Check(fs3_sync(m_h), "fs3_sync"); FetchParticles(); }
	}

	// ---- beyond the reference surface -------------------------------------------------------------
	void SetParticleReadback(bool every_step) { m_readback = every_step; }
	void FetchParticles() {
		m_particles.resize((size_t)fs3_num_particles(m_h));
		if (!m_particles.empty()) Check(fs3_get_particles(m_h, reinterpret_cast<fs3_particle*>(m_particles.data())), "fs3_get_particles");
	}
	// Call after editing m_particles on the host; the next Simulate() uploads the vector first.
	void MarkParticlesEdited() { m_dirty = true; }
	void Sync() { Check(fs3_sync(m_h), "fs3_sync"); }
	// private members of the reference class, exposed read-only (m_MU, m_MV, m_MW, m_Phi; Simulation3D.h:147-158)
	void GetVelocity(std::vector<float>& U, std::vector<float>& V, std::vector<float>& W) {
		U.resize((size_t)(mX + 1) * mY * mZ); V.resize((size_t)mX * (mY + 1) * mZ); W.resize((size_t)mX * mY * (mZ + 1));
		Check(fs3_get_velocity(m_h, U.data(), V.data(), W.data()), "fs3_get_velocity");
	}
	void GetPhi(std::vector<float>& phi) {
		phi.resize((size_t)mX * mY * mZ);
		Check(fs3_get_phi(m_h, phi.data()), "fs3_get_phi");
	}
	// on-disk state (fs3.h: lossless replacement of the reference's debugroutines.h text dumps)
	void SaveState(const std::string& path) { Check(fs3_save_state(m_h, path.c_str()), "fs3_save_state"); }
	void LoadState(const std::string& path) {
		Check(fs3_load_state(m_h, path.c_str()), "fs3_load_state");
		m_dirty = false;
		if (m_readback) FetchParticles();
	}
	// what GPFluidSim hands to its renderer (Simulation.h:118-140): phi and its blurred copy as device buffers
	const float* PhiDevice() { const float* p = nullptr; Check(fs3_phi_device(m_h, &p), "fs3_phi_device"); return p; }
	const float* BlurredPhiDevice() {
		const float* p = nullptr;
		Check(fs3_blur_phi(m_h), "fs3_blur_phi");
		Check(fs3_phi_blurred_device(m_h, &p), "fs3_phi_blurred_device");
		return p;
	}
	fs3_sim* Handle() { return m_h; }

	std::vector<Particle3> m_particles;   // creation order, as in the reference (never reordered)

private:
	void PushParticles() {
		Check(fs3_set_particles(m_h, reinterpret_cast<const fs3_particle*>(m_particles.data()), (long long)m_particles.size()), "fs3_set_particles");
		m_dirty = false;
	}
	void Check(int rc, const char* what) {
		if (rc != FS3_OK) throw std::runtime_error(std::string(what) + ": " + fs3_last_error(m_h));
	}
	fs3_sim* m_h = nullptr;
	int mX, mY, mZ;
	bool m_readback = true;
	bool m_dirty = false;
};
#endif
