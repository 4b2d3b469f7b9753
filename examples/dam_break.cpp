// dam_break.cpp -- headless C++ host driver over the drop-in class (include/FluidSim3Compat.hpp).
//
// Uses the solver exactly the way the reference's only caller does (FluidSimDemoOld3D.cpp): a FluidSim3
// object constructed as (n, n, n, (float)n) (:136), Simulate(dt) once per frame (:234), every
// m_particles[i].X/Y/Z read after the step (:263-269), ResetSimulation() on request (:437-439).  The window,
// vertex buffer and renderer of that file are replaced by a summary line per frame and an optional raw dump
// of m_particles (24 B per particle, creation order) for the parity tests.
//
//   g++ -std=c++17 -Iinclude examples/dam_break.cpp -Lfluidsimulation_b200 -lfluidsim3_b200 -o dam_break
//   ./dam_break [n=64] [frames=10] [dump.bin]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include "FluidSim3Compat.hpp"

int main(int argc, char** argv) {
	const int n = argc > 1 ? std::atoi(argv[1]) : 64;
	const int frames = argc > 2 ? std::atoi(argv[2]) : 10;
	const char* dump = argc > 3 ? argv[3] : nullptr;
	try {
		FluidSim3 fluidSim(n, n, n, (float)n);
		std::printf("grid %d^3, %zu particles\n", n, fluidSim.m_particles.size());
		for (int f = 0; f < frames; f++) {
			auto t0 = std::chrono::steady_clock::now();
			fluidSim.Simulate(0.01f);
			double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
			// per-frame consumer of the particle list: centre of mass and lowest/highest point of the water
			double cx = 0, cy = 0, cz = 0;
			float ymin = 1e30f, ymax = -1e30f;
			for (size_t i = 0; i < fluidSim.m_particles.size(); i++) {
				const Particle3& p = fluidSim.m_particles[i];
				cx += p.X; cy += p.Y; cz += p.Z;
				ymin = p.Y < ymin ? p.Y : ymin;
				ymax = p.Y > ymax ? p.Y : ymax;
			}
			const double inv = 1.0 / (double)fluidSim.m_particles.size();
			std::printf("frame %3d  %8.2f ms  centre (%.5f %.5f %.5f)  y in [%.5f, %.5f]\n", f + 1, ms, cx * inv, cy * inv, cz * inv, ymin, ymax);
		}
		if (dump) {
			FILE* fp = std::fopen(dump, "wb");
			if (!fp) { std::perror(dump); return 2; }
			std::fwrite(fluidSim.m_particles.data(), sizeof(Particle3), fluidSim.m_particles.size(), fp);
			std::fclose(fp);
		}
	} catch (const std::exception& e) {
		std::fprintf(stderr, "error: %s\n", e.what());
		return 1;
	}
	return 0;
}
