/* TEST INFRASTRUCTURE ONLY -- CPU oracle for the phi / binned-particle consumer interface.
 *
 * PARITY UNPINNED: the functions below restate arithmetic that exists in the reference only as D3D11 shaders
 * (FX/gpBlur.hlsl, FX/gpBinParticles.hlsl, FX/Render.fx), which cannot be run anywhere in this project (no graphics
 * stack: "unavailable on this box"), and the reference holds no golden vectors for them.  Two of the three are not even
 * deterministic in the reference (in-place racy blur, InterlockedAdd arrival order) and the third uses the hardware's
 * fixed-point texture filter.  The restatements take the deterministic reading stated next to each function; the CUDA
 * kernels in fluidsimulation_b200/csrc/consumer.cu are held bit-exact to THESE.  Plain C99, -ffp-contract=off.
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

/* FX/gpBlur.hlsl:6-17 -- grid[c] = (grid[c] + grid[c+x] + grid[c-x] + grid[c+y] + grid[c-y] + grid[c+z] + grid[c-z]) / 7.0f.
 * Out-of-range texture reads return 0 (D3D11; DTid - 1 wraps to a huge uint at 0).  The shader writes in place while
 * other threads read (racy); reading: every tap sees the input (Jacobi). */
void fs3o_blur_phi(const float* phi, float* out, int nx, int ny, int nz) {
	size_t plane = (size_t)nx * ny;
	for (int z = 0; z < nz; z++)
		for (int y = 0; y < ny; y++)
			for (int x = 0; x < nx; x++) {
				size_t q = (size_t)x + (size_t)nx * y + plane * z;
				float a = phi[q];
				a = a + (x + 1 < nx ? phi[q + 1] : 0.0f);
				a = a + (x > 0 ? phi[q - 1] : 0.0f);
				a = a + (y + 1 < ny ? phi[q + nx] : 0.0f);
				a = a + (y > 0 ? phi[q - nx] : 0.0f);
				a = a + (z + 1 < nz ? phi[q + plane] : 0.0f);
				a = a + (z > 0 ? phi[q - plane] : 0.0f);
				out[q] = a / 7.0f;
			}
}

/* One axis of Render.fx:77-115 map(): mp = mM*p + 0.5 (:81); i = floor(mp); f = mp - i (:82-83); on x and z
 * f = f*f*f*(f*(f*6 - 15) + 10) (:85); mp = i + f (:111); p = invmM*(mp - 0.5) (:113); then the linear sampler
 * (Render.fx:35-41: MIN_MAG_LINEAR, clamp addressing) at normalised coordinate p: texel space t = p*n - 0.5, taps
 * floor(t) and floor(t)+1 clamped to [0, n-1], weight t - floor(t).  The hardware filter's fixed-point weights are
 * replaced by the fp32 value. */
static void map_axis(float p, float n, float inv, int last, int smooth, int* i0, int* i1, float* w) {
	float mp = n * p + 0.5f;
	float i = floorf(mp);
	float f = mp - i;
	if (smooth) {
		float a = (f * f) * f;
		float b = f * (f * 6.0f - 15.0f) + 10.0f;
		f = a * b;
	}
	mp = i + f;
	float u = inv * (mp - 0.5f);
	float t = u * n - 0.5f;
	float fl = floorf(t);
	*w = t - fl;
	if (fl < -1.0f) fl = -1.0f;
	if (fl > (float)last) fl = (float)last;
	int k = (int)fl;
	*i0 = k < 0 ? 0 : k;
	*i1 = k + 1 > last ? last : k + 1;
}
static float lerpf(float a, float b, float t) { return a + (b - a) * t; }
/* Render.fx:115: gPhi.SampleLevel(sam, p, 0) * invmM.x; filter order x, y, z. */
void fs3o_sample_phi(const float* phi, int nx, int ny, int nz, const float* xyz, long count, float* out) {
	float ix = 1.0f / (float)nx, iy = 1.0f / (float)ny, iz = 1.0f / (float)nz;   /* Render.fx:531-533 */
	size_t pl = (size_t)nx * ny;
	for (long i = 0; i < count; i++) {
		int x0, x1, y0, y1, z0, z1;
		float wx, wy, wz;
		map_axis(xyz[3 * i + 0], (float)nx, ix, nx - 1, 1, &x0, &x1, &wx);
		map_axis(xyz[3 * i + 1], (float)ny, iy, ny - 1, 0, &y0, &y1, &wy);
		map_axis(xyz[3 * i + 2], (float)nz, iz, nz - 1, 1, &z0, &z1, &wz);
		const float* r00 = phi + (size_t)nx * y0 + pl * z0;
		const float* r10 = phi + (size_t)nx * y1 + pl * z0;
		const float* r01 = phi + (size_t)nx * y0 + pl * z1;
		const float* r11 = phi + (size_t)nx * y1 + pl * z1;
		float a00 = lerpf(r00[x0], r00[x1], wx), a10 = lerpf(r10[x0], r10[x1], wx);
		float a01 = lerpf(r01[x0], r01[x1], wx), a11 = lerpf(r11[x0], r11[x1], wx);
		out[i] = lerpf(lerpf(a00, a10, wy), lerpf(a01, a11, wy), wz) * ix;
	}
}

/* FX/gpCountParticles.hlsl:11-26 + host prefix sum Simulation.cpp:660-686 + FX/gpBinParticles.hlsl:13-31, with the CPU
 * path's cell rule (cell = (int)roundf(n*X), Simulation3D.cpp:286-288) and a deterministic order inside a cell:
 * by octant (z half, y half, x half; upper half = n*X >= cell centre), then creation index. */
void fs3o_bin_particles(const float* part6, long np, int nx, int ny, int nz, float cpm, float* binned6, unsigned* offsets) {
	size_t nc = (size_t)nx * ny * nz;
	unsigned* cursor = (unsigned*)calloc(8 * nc + 1, sizeof(unsigned));
	long* key = (long*)malloc(sizeof(long) * (size_t)(np > 0 ? np : 1));
	for (long i = 0; i < np; i++) {
		float q[3] = {part6[6 * i] * cpm, part6[6 * i + 1] * cpm, part6[6 * i + 2] * cpm};
		int c[3], o[3];
		for (int a = 0; a < 3; a++) { float r = roundf(q[a]); c[a] = (int)r; o[a] = q[a] >= r ? 1 : 0; }
		if (c[0] < 0 || c[0] >= nx || c[1] < 0 || c[1] >= ny || c[2] < 0 || c[2] >= nz) { key[i] = -1; continue; }
		size_t cell = (size_t)c[0] + (size_t)nx * (c[1] + (size_t)ny * c[2]);
		key[i] = (long)(8 * cell + (size_t)(o[0] + 2 * o[1] + 4 * o[2]));
		cursor[key[i] + 1]++;
	}
	for (size_t k = 0; k < 8 * nc; k++) cursor[k + 1] += cursor[k];
	for (size_t c = 0; c <= nc; c++) offsets[c] = cursor[8 * c];
	for (long i = 0; i < np; i++) {
		if (key[i] < 0) continue;
		unsigned slot = cursor[key[i]]++;
		memcpy(binned6 + 6 * (size_t)slot, part6 + 6 * i, 24);
	}
	free(key);
	free(cursor);
}
