// On-disk simulation state: one binary file holding everything Simulate() reads at the
// start of a step -- the particles (creation order, reference wire layout) and the MAC grids -- plus phi and the
// parameters, so that a run can be stopped and resumed bit-identically.  It replaces the reference's text dumps
// (debugroutines.h:24-178: exportArray / ReadArrayFromFile write tab-separated decimal text, which is lossy -- the
// author's own "stretch goal: save floats in hexfloat format for lossless passing", debugroutines.h:28-29 -- and has
// no particle import).  Pure host code above the C ABI's own get/set calls; nothing here touches a kernel.
//
// Layout (little-endian, no padding between sections):
//     header (256 bytes, fs3_state_disk_header) | particles (num_particles x 24 B) | [creation ids (num_particles x 4 B)]
//     | U | V | W | phi   (fp32, reference layout, Simulation3D.h:41-52)
// The id section exists only in a per-rank file of a z-slab sharded run (one file per rank, each with the rank's own
// particles and the whole grids, which every rank holds).  `checksum` covers everything after the header.
#include "fs3_internal.cuh"
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <new>

void fs3i_set_global_error(const std::string& msg);   // fs3_api.cu: the message fs3_last_error(NULL) returns

#define ENTER(s)                                                       \
	if (!(s)) return FS3_ERR_INVALID;                                  \
	do {                                                               \
		cudaError_t e0_ = cudaSetDevice((s)->device);                  \
		if (e0_ != cudaSuccess) { (s)->err = cudaGetErrorString(e0_); return FS3_ERR_CUDA; } \
	} while (0)

namespace {

const char kMagic[8] = {'F', 'S', '3', 'S', 'T', 'A', 'T', 'E'};
const uint32_t kVersion = 1;

struct fs3_state_disk_header {
	char magic[8];
	uint32_t version, header_bytes;
	int32_t nx, ny, nz;
	float cells_per_meter;
	int64_t num_particles, total_particles, steps;
	int32_t rank, world, has_ids, precision;
	int32_t sor_iterations, force_dt_0_01;
	double omega;
	float rho, gravity_y, nu, particle_radius;
	int64_t np_all[FS3_MAX_WORLD];
	uint64_t payload_bytes, checksum;
	int32_t levelset_mode;   // 0 in files written before the field existed (the reserved word was zero)
	int32_t extrapolation_mode;
};
static_assert(sizeof(fs3_state_disk_header) == 256, "state file header must be 256 bytes");

struct Sizes {
	size_t nU, nV, nW, nC;
};
bool grid_sizes(int nx, int ny, int nz, Sizes* z) {
	if (nx < 1 || ny < 1 || nz < 1 || nx > 4096 || ny > 4096 || nz > 4096) return false;
	z->nU = (size_t)(nx + 1) * ny * nz; z->nV = (size_t)nx * (ny + 1) * nz; z->nW = (size_t)nx * ny * (nz + 1); z->nC = (size_t)nx * ny * nz;
	return true;
}
uint64_t payload_size(const fs3_state_info* in, const Sizes& z) {
	return (uint64_t)in->num_particles * (sizeof(fs3_particle) + (in->has_ids ? 4u : 0u)) + 4ull * (z.nU + z.nV + z.nW + z.nC);
}

int io_fail(const std::string& msg) { fs3i_set_global_error(msg); return FS3_ERR_IO; }
int bad(const std::string& msg) { fs3i_set_global_error(msg); return FS3_ERR_INVALID; }

void to_disk(const fs3_state_info* in, fs3_state_disk_header* h) {
	memset(h, 0, sizeof(*h));
	memcpy(h->magic, kMagic, 8);
	h->version = kVersion; h->header_bytes = sizeof(*h);
	h->nx = in->nx; h->ny = in->ny; h->nz = in->nz; h->cells_per_meter = in->cells_per_meter;
	h->num_particles = in->num_particles; h->total_particles = in->total_particles; h->steps = in->steps;
	h->rank = in->rank; h->world = in->world; h->has_ids = in->has_ids ? 1 : 0;
	h->precision = in->params.precision; h->sor_iterations = in->params.sor_iterations; h->force_dt_0_01 = in->params.force_dt_0_01;
	h->omega = in->params.omega; h->rho = in->params.rho; h->gravity_y = in->params.gravity_y; h->nu = in->params.nu;
	h->particle_radius = in->params.particle_radius; h->levelset_mode = in->params.levelset_mode;
	h->extrapolation_mode = in->params.extrapolation_mode;
	for (int r = 0; r < FS3_MAX_WORLD; r++) h->np_all[r] = in->rank_particles[r];
}
void from_disk(const fs3_state_disk_header* h, fs3_state_info* out) {
	memset(out, 0, sizeof(*out));
	out->version = (int)h->version;
	out->nx = h->nx; out->ny = h->ny; out->nz = h->nz; out->cells_per_meter = h->cells_per_meter;
	out->num_particles = h->num_particles; out->total_particles = h->total_particles; out->steps = h->steps;
	out->rank = h->rank; out->world = h->world; out->has_ids = h->has_ids;
	out->params.precision = h->precision; out->params.sor_iterations = h->sor_iterations; out->params.force_dt_0_01 = h->force_dt_0_01;
	out->params.omega = h->omega; out->params.rho = h->rho; out->params.gravity_y = h->gravity_y; out->params.nu = h->nu;
	out->params.particle_radius = h->particle_radius; out->params.levelset_mode = h->levelset_mode;
	out->params.extrapolation_mode = h->extrapolation_mode;
	for (int r = 0; r < FS3_MAX_WORLD; r++) out->rank_particles[r] = h->np_all[r];
	out->payload_bytes = h->payload_bytes; out->checksum = h->checksum;
}

// Reads and validates the header; leaves the file positioned at the payload.
int read_header(FILE* f, const char* path, fs3_state_disk_header* h, Sizes* z) {
	if (fread(h, 1, sizeof(*h), f) != sizeof(*h)) return io_fail(std::string(path) + ": shorter than a state-file header");
	if (memcmp(h->magic, kMagic, 8)) return io_fail(std::string(path) + ": not a FluidSim3 state file (bad magic)");
	if (h->version != kVersion || h->header_bytes != sizeof(*h)) return io_fail(std::string(path) + ": unsupported state-file version");
	if (!grid_sizes(h->nx, h->ny, h->nz, z) || h->num_particles < 0 || h->num_particles > 0x7fffffffLL || h->world < 1 || h->world > FS3_MAX_WORLD ||
	    h->rank < 0 || h->rank >= h->world || (h->has_ids != 0 && h->has_ids != 1))
		return io_fail(std::string(path) + ": corrupt state-file header");
	fs3_state_info tmp;
	from_disk(h, &tmp);
	if (h->payload_bytes != payload_size(&tmp, *z)) return io_fail(std::string(path) + ": corrupt state-file header (payload size does not match the shape)");
	return FS3_OK;
}

struct Section {
	void* ptr;
	size_t bytes;
};

}  // namespace

extern "C" {

// 64-bit FNV-1a over little-endian 8-byte words (the tail is zero-padded); chainable through `seed`.
unsigned long long fs3_state_checksum(const void* data, size_t bytes, unsigned long long seed) {
	const unsigned char* p = (const unsigned char*)data;
	uint64_t h = seed ? seed : 0xcbf29ce484222325ull;
	size_t nw = bytes / 8;
	for (size_t i = 0; i < nw; i++) {
		uint64_t w;
		memcpy(&w, p + 8 * i, 8);
		h = (h ^ w) * 0x100000001b3ull;
	}
	if (bytes & 7) {
		uint64_t w = 0;
		memcpy(&w, p + 8 * nw, bytes & 7);
		h = (h ^ w) * 0x100000001b3ull;
	}
	return h;
}

int fs3_state_file_info(const char* path, fs3_state_info* info) {
	if (!path || !info) return FS3_ERR_INVALID;
	FILE* f = fopen(path, "rb");
	if (!f) return io_fail(std::string(path) + ": cannot open for reading");
	fs3_state_disk_header h;
	Sizes z;
	int rc = read_header(f, path, &h, &z);
	fclose(f);
	if (rc) return rc;
	from_disk(&h, info);
	return FS3_OK;
}

// The checksum is chained over the sections in file order (each section's result seeds the next); a section whose
// size is not a multiple of 8 is hashed with a zero-padded tail word, which the reader reproduces section by section.
int fs3_state_file_write(const char* path, const fs3_state_info* info, const fs3_particle* particles, const int* ids,
                         const float* U, const float* V, const float* W, const float* phi) {
	if (!path || !info || !U || !V || !W || !phi) return FS3_ERR_INVALID;
	Sizes z;
	if (!grid_sizes(info->nx, info->ny, info->nz, &z)) return bad("state file: bad grid shape");
	if (info->num_particles < 0 || info->num_particles > 0x7fffffffLL || (info->num_particles > 0 && !particles)) return bad("state file: bad particle count");
	if (info->has_ids && info->num_particles > 0 && !ids) return bad("state file: has_ids set but no id array given");
	if (info->world < 1 || info->world > FS3_MAX_WORLD || info->rank < 0 || info->rank >= info->world) return bad("state file: bad rank / world");
	fs3_state_disk_header h;
	to_disk(info, &h);
	const size_t np = (size_t)info->num_particles;
	Section sec[6] = {{(void*)particles, np * sizeof(fs3_particle)}, {(void*)ids, info->has_ids ? np * 4 : 0}, {(void*)U, z.nU * 4}, {(void*)V, z.nV * 4},
	                  {(void*)W, z.nW * 4}, {(void*)phi, z.nC * 4}};
	h.payload_bytes = payload_size(info, z);
	uint64_t sum = 0;
	for (const Section& s : sec) if (s.bytes) sum = fs3_state_checksum(s.ptr, s.bytes, sum);
	h.checksum = sum;
	// write beside the target and rename: a crash never leaves a truncated file under the final name
	const std::string tmp = std::string(path) + ".tmp";
	FILE* f = fopen(tmp.c_str(), "wb");
	if (!f) return io_fail(tmp + ": cannot open for writing");
	bool ok = fwrite(&h, 1, sizeof(h), f) == sizeof(h);
	for (const Section& s : sec) if (ok && s.bytes) ok = fwrite(s.ptr, 1, s.bytes, f) == s.bytes;
	ok = (fclose(f) == 0) && ok;
	if (!ok) { remove(tmp.c_str()); return io_fail(tmp + ": write failed (disk full?)"); }
	if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return io_fail(std::string(path) + ": cannot rename the finished file into place"); }
	return FS3_OK;
}

// Buffers are caller-owned and sized from fs3_state_file_info; any of them may be NULL to skip that section (the
// checksum is then not verified, because it needs every byte).
int fs3_state_file_read(const char* path, fs3_state_info* info, fs3_particle* particles, int* ids, float* U, float* V, float* W, float* phi) {
	if (!path) return FS3_ERR_INVALID;
	FILE* f = fopen(path, "rb");
	if (!f) return io_fail(std::string(path) + ": cannot open for reading");
	fs3_state_disk_header h;
	Sizes z;
	int rc = read_header(f, path, &h, &z);
	if (rc) { fclose(f); return rc; }
	const size_t np = (size_t)h.num_particles;
	Section sec[6] = {{particles, np * sizeof(fs3_particle)}, {ids, h.has_ids ? np * 4 : 0}, {U, z.nU * 4}, {V, z.nV * 4}, {W, z.nW * 4}, {phi, z.nC * 4}};
	bool all = true;
	uint64_t sum = 0;
	for (const Section& s : sec) {
		if (!s.bytes) continue;
		if (!s.ptr) {
			all = false;
			if (fseek(f, (long)s.bytes, SEEK_CUR) != 0) { fclose(f); return io_fail(std::string(path) + ": seek failed"); }
			continue;
		}
		if (fread(s.ptr, 1, s.bytes, f) != s.bytes) { fclose(f); return io_fail(std::string(path) + ": truncated state file"); }
		sum = fs3_state_checksum(s.ptr, s.bytes, sum);
	}
	fclose(f);
	if (all && sum != h.checksum) return io_fail(std::string(path) + ": checksum mismatch (the payload is corrupt)");
	if (info) from_disk(&h, info);
	return FS3_OK;
}

static int fail_from_global(fs3_sim* s, int rc) { s->err = fs3_last_error(nullptr); return rc; }

// Writes the handle's current state.  Sharded handle: collective in the sense that every rank calls it, each with
// its OWN path; the file holds the rank's particles with their creation ids and the whole grids.
int fs3_save_state(fs3_sim* s, const char* path) {
	ENTER(s);
	if (!path) return FS3_ERR_INVALID;
	fs3_state_info in;
	memset(&in, 0, sizeof(in));
	in.version = (int)kVersion; in.nx = s->nx; in.ny = s->ny; in.nz = s->nz; in.cells_per_meter = s->cpm;
	in.total_particles = fs3_num_particles(s); in.steps = s->steps; in.rank = s->rank; in.world = s->world; in.has_ids = s->world > 1;
	in.params = s->prm;
	for (int r = 0; r < s->world; r++) in.rank_particles[r] = s->world > 1 ? s->np_all[r] : s->np;
	std::vector<fs3_particle> part;
	std::vector<int> ids;
	std::vector<float> U, V, W, phi;
	try {
		U.resize(s->nU); V.resize(s->nV); W.resize(s->nW); phi.resize(s->nC);
		int rc;
		if (s->world > 1) {
			long long cnt = 0;
			part.resize((size_t)std::max<long long>(s->np, 1)); ids.resize(part.size());
			if ((rc = fs3_get_local_particles(s, part.data(), ids.data(), (long long)part.size(), &cnt))) return rc;
			in.num_particles = cnt;
		} else {
			in.num_particles = s->np;
			part.resize((size_t)std::max<long long>(s->np, 1));
			if (s->np > 0 && (rc = fs3_get_particles(s, part.data()))) return rc;
		}
		if ((rc = fs3_get_velocity(s, U.data(), V.data(), W.data()))) return rc;
		if ((rc = fs3_get_phi(s, phi.data()))) return rc;
	} catch (const std::bad_alloc&) {
		s->err = "out of host memory while staging the state";
		return FS3_ERR_INVALID;
	}
	int rc = fs3_state_file_write(path, &in, part.data(), in.has_ids ? ids.data() : nullptr, U.data(), V.data(), W.data(), phi.data());
	return rc ? fail_from_global(s, rc) : FS3_OK;
}

// Restores a state written by fs3_save_state into a handle of the same shape (and, sharded, the same rank / world):
// particles, grids, phi, the solver parameters and the step counter.  The next fs3_step continues bit-identically.
int fs3_load_state(fs3_sim* s, const char* path) {
	ENTER(s);
	if (!path) return FS3_ERR_INVALID;
	fs3_state_info in;
	int rc = fs3_state_file_info(path, &in);
	if (rc) return fail_from_global(s, rc);
	if (in.nx != s->nx || in.ny != s->ny || in.nz != s->nz || in.cells_per_meter != s->cpm) { s->err = "state file is for a different grid"; return FS3_ERR_INVALID; }
	if (in.world != s->world || in.rank != s->rank || (in.has_ids != 0) != (s->world > 1)) { s->err = "state file was written by a different rank / world size"; return FS3_ERR_INVALID; }
	if (s->world > 1 && in.num_particles > s->cap) { s->err = "state file holds more particles than the slab's capacity"; return FS3_ERR_STATE; }
	std::vector<fs3_particle> part;
	std::vector<int> ids;
	std::vector<float> U, V, W, phi;
	try {
		part.resize((size_t)std::max<long long>(in.num_particles, 1));
		if (in.has_ids) ids.resize(part.size());
		U.resize(s->nU); V.resize(s->nV); W.resize(s->nW); phi.resize(s->nC);
	} catch (const std::bad_alloc&) {
		s->err = "out of host memory while staging the state";
		return FS3_ERR_INVALID;
	}
	if ((rc = fs3_state_file_read(path, &in, part.data(), in.has_ids ? ids.data() : nullptr, U.data(), V.data(), W.data(), phi.data()))) return fail_from_global(s, rc);
	// parameters first: a change of precision re-allocates the pressure system
	fs3_params p = in.params;
	p.record_residuals = s->prm.record_residuals; p.record_stage_times = s->prm.record_stage_times;
	if ((rc = fs3_set_params(s, &p))) return rc;
	if (s->world > 1) {
		if ((rc = fs3_set_local_particles(s, part.data(), ids.data(), in.num_particles))) return rc;
		for (int r = 0; r < s->world; r++) s->np_all[r] = in.rank_particles[r];   // the other ranks load their own files
	} else {
		if ((rc = fs3_set_particles(s, part.data(), in.num_particles))) return rc;
	}
	if ((rc = fs3_set_velocity(s, U.data(), V.data(), W.data()))) return rc;
	if ((rc = fs3_set_phi(s, phi.data()))) return rc;
	s->steps = in.steps;
	return FS3_OK;
}

long long fs3_step_count(const fs3_sim* s) { return s ? s->steps : 0; }

}  // extern "C"
