/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the "line-sweep" level set: the reference's GPU
 * formulation of ComputeLevelSet -- clear phi to +INF (Simulation.cpp:718-723), closest particle among the 27 neighbour
 * cells (FX/gpComputeClosestParticleNeighbors.hlsl:40-112), then 24 one-directional line sweeps
 * (FX/gpClosestParticlesSweep{Xm,Xp,Ym,Yp,Zm,Zp}.hlsl, order table Simulation.cpp:744-753).
 *
 * PARITY UNPINNED: these exist in the reference only as D3D11 compute shaders, which cannot be run anywhere in this project
 * (no graphics stack: "unavailable on this box"), and the reference holds no golden vectors for them.  Where the shaders'
 * result depends on something the reference leaves unspecified, the reading below is taken and the CUDA kernels
 * (fluidsimulation_b200/csrc/levelset_lines.cu) are held bit-exact to it:
 *   * particle -> cell: the solver's own rule (int)roundf(n*X) (Simulation3D.cpp:286-288), not the shader's
 *     uint3(mM*p + 0.5f) (FX/gpCountParticles.hlsl:22): the stage consumes the step's own cell lists;
 *   * order of the particles inside a cell (InterlockedAdd arrival in FX/gpBinParticles.hlsl:24-29): by octant, then
 *     creation index -- it only matters for exact distance ties (the first candidate scanned wins, strict `<`);
 *   * gClosestParticles is never cleared between frames (no clear dispatch in Simulation.cpp:718-733) and starts as zeros:
 *     every pass starts from the FIRST frame's state, i.e. every cell points at binned particle 0 (the first particle of
 *     the lowest-numbered non-empty cell);
 *   * distance(a, b) = sqrtf(dx*dx + dy*dy + dz*dz), products summed left to right, never fused.
 * Quirks that ARE specified by the shader source and are kept: the `+` sweeps take their first candidate from the cell at
 * coordinate 0 of the line, not from the far end where they start (FX/gpClosestParticlesSweepXp.hlsl:22-28), and visit
 * coordinate 0 itself; the `-` sweeps start at coordinate 1.  Plain C99, -ffp-contract=off.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

static float gls_dist(const float* q, float cx, float cy, float cz, float radius) {
	float dx = cx - q[0], dy = cy - q[1], dz = cz - q[2];
	return sqrtf(dx * dx + dy * dy + dz * dz) - radius;
}

/* part6: AoS X,Y,Z,uX,uY,uZ in creation order.  Out: phi[n^3], closest[n^3] (creation index).  Returns 0, or -1 when no
 * particle lies inside the grid (the shaders would read gParticles[0] of an empty buffer). */
int fs3o_gls_levelset(const float* part6, long np, int nx, int ny, int nz, float cpm, float radius, float* phi, int* closest) {
	size_t nc = (size_t)nx * ny * nz;
	/* --- bin: cell-major, octant, creation index (FX/gpCountParticles + host prefix sum + FX/gpBinParticles) --- */
	unsigned* cursor = (unsigned*)calloc(8 * nc + 1, sizeof(unsigned));
	long* key = (long*)malloc(sizeof(long) * (size_t)(np > 0 ? np : 1));
	long inside = 0;
	for (long i = 0; i < np; i++) {
		int c[3], o[3];
		for (int a = 0; a < 3; a++) { float q = part6[6 * i + a] * cpm; float r = roundf(q); c[a] = (int)r; o[a] = q >= r ? 1 : 0; }
		if (c[0] < 0 || c[0] >= nx || c[1] < 0 || c[1] >= ny || c[2] < 0 || c[2] >= nz) { key[i] = -1; continue; }
		key[i] = (long)(8 * ((size_t)c[0] + (size_t)nx * (c[1] + (size_t)ny * c[2])) + (size_t)(o[0] + 2 * o[1] + 4 * o[2]));
		cursor[key[i] + 1]++;
		inside++;
	}
	if (inside == 0) { free(cursor); free(key); return -1; }
	for (size_t k = 0; k < 8 * nc; k++) cursor[k + 1] += cursor[k];
	unsigned* offsets = (unsigned*)malloc(sizeof(unsigned) * (nc + 1));
	for (size_t c = 0; c <= nc; c++) offsets[c] = cursor[8 * c];
	int* order = (int*)malloc(sizeof(int) * (size_t)inside);          /* binned slot -> creation index */
	for (long i = 0; i < np; i++) if (key[i] >= 0) order[cursor[key[i]]++] = (int)i;
	float* q = (float*)malloc(sizeof(float) * 3 * (size_t)np);          /* mM * pos, as the shaders compute it */
	for (long i = 0; i < np; i++) for (int a = 0; a < 3; a++) q[3 * i + a] = part6[6 * i + a] * cpm;

	/* --- clear: phi = +INF (Simulation.cpp:718-723); closest = binned particle 0 --- */
	for (size_t c = 0; c < nc; c++) { phi[c] = INFINITY; closest[c] = order[0]; }

	/* --- closest particle among the 27 neighbour cells, z outermost, x innermost (gpComputeClosestParticleNeighbors.hlsl:89-109);
	 *     out-of-range neighbours hold no particles (their offset/size pair wraps to an empty range, gpParticleIndexing.hlsli:28-33) --- */
	for (int z = 0; z < nz; z++)
		for (int y = 0; y < ny; y++)
			for (int x = 0; x < nx; x++) {
				size_t c = (size_t)x + (size_t)nx * (y + (size_t)ny * z);
				for (int dz = -1; dz <= 1; dz++)
					for (int dy = -1; dy <= 1; dy++)
						for (int dx = -1; dx <= 1; dx++) {
							int X = x + dx, Y = y + dy, Z = z + dz;
							if (X < 0 || X >= nx || Y < 0 || Y >= ny || Z < 0 || Z >= nz) continue;
							size_t o = (size_t)X + (size_t)nx * (Y + (size_t)ny * Z);
							for (unsigned i = offsets[o]; i < offsets[o + 1]; i++) {
								float d = gls_dist(q + 3 * order[i], (float)x, (float)y, (float)z, radius);
								if (d < phi[c]) { phi[c] = d; closest[c] = order[i]; }
							}
						}
			}

	/* --- 24 line sweeps (Simulation.cpp:744-794).  0 1 2 3 4 5 = x- x+ y- y+ z- z+ --- */
	static const int dirs[24] = {0, 2, 4, 1, 2, 4, 0, 3, 4, 1, 3, 4, 0, 2, 5, 1, 2, 5, 0, 3, 5, 1, 3, 5};
	const int dims[3] = {nx, ny, nz};
	const size_t strides[3] = {1, (size_t)nx, (size_t)nx * ny};
	for (int s = 0; s < 24; s++) {
		const int axis = dirs[s] >> 1, plus = dirs[s] & 1;
		const int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3, len = dims[axis];
		for (int v = 0; v < dims[a2]; v++)
			for (int u = 0; u < dims[a1]; u++) {
				int pos[3];
				pos[a1] = u; pos[a2] = v; pos[axis] = 0;
				const size_t line = (size_t)pos[0] * strides[0] + (size_t)pos[1] * strides[1] + (size_t)pos[2] * strides[2];
				int last = closest[line];                                 /* the cell at coordinate 0, in both directions */
				for (int k = plus ? len - 1 : 1; plus ? k >= 0 : k < len; k += plus ? -1 : 1) {
					const size_t c = line + (size_t)k * strides[axis];
					pos[axis] = k;
					float d = gls_dist(q + 3 * last, (float)pos[0], (float)pos[1], (float)pos[2], radius);
					if (d < phi[c]) { phi[c] = d; closest[c] = last; }
					else last = closest[c];
				}
			}
	}
	free(q); free(order); free(offsets); free(key); free(cursor);
	return 0;
}

/* FX/gpExtrapolateParticleVelocities.hlsl:16-42 -- one layer: an invalid node (the shader's +INF marker; here valid[q] == 0)
 * becomes the mean of its valid 6-neighbours, summed in the order +x -x +y -y +z -z and divided by their number, or 0 when
 * it has none.  Out-of-range neighbours are texture reads that return 0 (finite), so they count as valid zeros.  The shader
 * updates the grid in place while neighbours read it (racy); reading: only nodes valid before the pass are neighbours. */
void fs3o_gls_extrapolate(float* grid, const unsigned char* valid, int xs, int ys, int zs) {
	size_t plane = (size_t)xs * ys, n = plane * zs;
	float* out = (float*)malloc(sizeof(float) * n);
	memcpy(out, grid, sizeof(float) * n);
	for (int z = 0; z < zs; z++)
		for (int y = 0; y < ys; y++)
			for (int x = 0; x < xs; x++) {
				size_t q = (size_t)x + (size_t)xs * y + plane * z;
				if (valid[q]) continue;
				float sum = 0.0f, num = 0.0f;
				const int inside[6] = {x + 1 < xs, x > 0, y + 1 < ys, y > 0, z + 1 < zs, z > 0};
				const size_t off[6] = {q + 1, q - 1, q + xs, q - xs, q + plane, q - plane};
				for (int k = 0; k < 6; k++) {
					if (!inside[k]) num = num + 1.0f;
					else if (valid[off[k]]) { sum = sum + grid[off[k]]; num = num + 1.0f; }
				}
				out[q] = num != 0.0f ? sum / num : 0.0f;
			}
	memcpy(grid, out, sizeof(float) * n);
	free(out);
}
