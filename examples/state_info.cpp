// state_info.cpp -- host-only C++ tool over the state-file part of the C ABI (include/fs3.h): prints the header of a
// FluidSim3 state file, verifies its checksum and, like the reference's debug helpers (debugroutines.h: ComputeL2Norm :152-162,
// exportArrayYSlice :68-89), reports the L2 norm of every array and can export one y-slice of a grid as tab-separated text --
// written with %.9g, which round-trips fp32 (the reference prints 6 significant digits and loses bits, :34).  No GPU needed.
//
//   g++ -std=c++17 -Iinclude examples/state_info.cpp -Lfluidsimulation_b200 -lfluidsim3_b200 -o state_info
//   ./state_info file.fs3 [phi|U|V|W yLayer slice.tsv]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "fs3.h"

static double l2(const float* a, size_t n) {   // debugroutines.h:152-162
	double s = 0.0;
	for (size_t i = 0; i < n; i++) s += (double)a[i] * (double)a[i];
	return std::sqrt(s);
}

int main(int argc, char** argv) {
	if (argc != 2 && argc != 5) { std::fprintf(stderr, "usage: %s file.fs3 [phi|U|V|W yLayer slice.tsv]\n", argv[0]); return 2; }
	fs3_state_info in;
	if (fs3_state_file_info(argv[1], &in) != FS3_OK) { std::fprintf(stderr, "error: %s\n", fs3_last_error(nullptr)); return 1; }
	const size_t nx = (size_t)in.nx, ny = (size_t)in.ny, nz = (size_t)in.nz, np = (size_t)in.num_particles;
	std::vector<fs3_particle> part(np ? np : 1);
	std::vector<int> ids(in.has_ids ? (np ? np : 1) : 0);
	std::vector<float> U((nx + 1) * ny * nz), V(nx * (ny + 1) * nz), W(nx * ny * (nz + 1)), phi(nx * ny * nz);
	if (fs3_state_file_read(argv[1], &in, part.data(), in.has_ids ? ids.data() : nullptr, U.data(), V.data(), W.data(), phi.data()) != FS3_OK) {
		std::fprintf(stderr, "error: %s\n", fs3_last_error(nullptr));
		return 1;
	}
	std::printf("version %d  grid %dx%dx%d  cells_per_meter %g  steps %lld\n", in.version, in.nx, in.ny, in.nz, (double)in.cells_per_meter, in.steps);
	std::printf("particles %lld of %lld  rank %d/%d  ids %s\n", in.num_particles, in.total_particles, in.rank, in.world, in.has_ids ? "yes" : "no");
	std::printf("params: sor_iterations %d  omega %g  precision %d  force_dt_0_01 %d  levelset_mode %d  rho %g  gravity_y %g  nu %g  radius %g\n",
	            in.params.sor_iterations, in.params.omega, in.params.precision, in.params.force_dt_0_01, in.params.levelset_mode,
	            (double)in.params.rho, (double)in.params.gravity_y, (double)in.params.nu, (double)in.params.particle_radius);
	std::printf("payload %llu bytes  checksum %016llx  OK\n", in.payload_bytes, in.checksum);
	std::printf("L2: particles %.9g  U %.9g  V %.9g  W %.9g  phi %.9g\n", l2(&part[0].X, 6 * np), l2(U.data(), U.size()), l2(V.data(), V.size()),
	            l2(W.data(), W.size()), l2(phi.data(), phi.size()));
	if (argc == 5) {
		const char* which = argv[2];
		const long yl = std::atol(argv[3]);
		const float* a = nullptr;
		size_t sx = 0, sy = 0, sz = 0;
		if (!std::strcmp(which, "phi")) { a = phi.data(); sx = nx; sy = ny; sz = nz; }
		else if (!std::strcmp(which, "U")) { a = U.data(); sx = nx + 1; sy = ny; sz = nz; }
		else if (!std::strcmp(which, "V")) { a = V.data(); sx = nx; sy = ny + 1; sz = nz; }
		else if (!std::strcmp(which, "W")) { a = W.data(); sx = nx; sy = ny; sz = nz + 1; }
		if (!a || yl < 0 || (size_t)yl >= sy) { std::fprintf(stderr, "error: no such array / y layer\n"); return 2; }
		FILE* fp = std::fopen(argv[4], "w");
		if (!fp) { std::perror(argv[4]); return 2; }
		for (size_t z = 0; z < sz; z++)          // one row per z, x across: the reference's exportArrayYSlice layout (:76-86)
			for (size_t x = 0; x < sx; x++) std::fprintf(fp, "%.9g%c", (double)a[x + sx * ((size_t)yl + sy * z)], x + 1 == sx ? (z + 1 == sz ? ' ' : '\n') : '\t');
		std::fclose(fp);
	}
	return 0;
}
