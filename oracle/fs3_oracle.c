/* TEST INFRASTRUCTURE ONLY -- CPU oracle (see fs3_oracle.h).  Plain C99.
 * Citations "S3D.cpp:a-b" / "S3D.h:a-b" are /root/reference/FluidSimulationDemo/Simulation3D.{cpp,h}.
 * All fp32 arithmetic keeps the reference's operation order; build with -ffp-contract=off.
 */
#include "fs3_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <float.h>

struct fs3o {
	int nx, ny, nz;
	float cpm;                      /* m_CellsPerMeter */
	float rho, gy, nu, radius;      /* S3D.h:135-145 */
	float *u, *v, *w, *phi;
	long np;
	float* part;                    /* AoS, 6 floats per particle (ParticleDefs.h:39-47) */
	/* retained locals (parity hooks) */
	int* closest;
	double *p, *b, *diag;
	int sor_iters;
	int record_res;
	double *res_l2, *res_linf;
	int n_res;
};

static size_t n_u(const fs3o* s) { return (size_t)(s->nx + 1) * s->ny * s->nz; }
static size_t n_v(const fs3o* s) { return (size_t)s->nx * (s->ny + 1) * s->nz; }
static size_t n_w(const fs3o* s) { return (size_t)s->nx * s->ny * (s->nz + 1); }
static size_t n_c(const fs3o* s) { return (size_t)s->nx * s->ny * s->nz; }

/* index maps, S3D.h:41-52 */
#define IU(s, i, j, k) ((i) + ((s)->nx + 1) * ((j) + (s)->ny * (k)))
#define IV(s, i, j, k) ((i) + (s)->nx * ((j) + ((s)->ny + 1) * (k)))
#define IW(s, i, j, k) ((i) + (s)->nx * ((j) + (s)->ny * (k)))
#define IC(s, i, j, k) ((i) + (s)->nx * ((j) + (s)->ny * (k)))

static float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }   /* MathHelper.h:49-53 */
static double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }
static float lerpf(float a, float b, float t) { return a + (b - a) * t; }                       /* MathHelper.h:43-47 */

/* ---- RNG: std::default_random_engine(0) + uniform_real_distribution<float>(-0.25f,0.25f)
 * as libstdc++ implements them (S3D.cpp:43-44).  minstd_rand0: x <- 16807 x mod (2^31-1), seed 0 -> 1.
 * generate_canonical<float,24> takes one draw: float(x - 1) / float(2147483646) where the divisor
 * rounds to 2^31 in fp32; results >= 1 are replaced by nextafterf(1,0). */
typedef struct { uint64_t x; } lcg_t;
static float lcg_draw(lcg_t* g) {
	g->x = (g->x * 16807ull) % 2147483647ull;
	float r = (float)(uint32_t)(g->x - 1) / 2147483648.0f;
	if (r >= 1.0f) r = nextafterf(1.0f, 0.0f);
	return r * 0.5f + (-0.25f);
}

/* one axis of InterpolateMACCell: clamp, floor, step back from the last node, fraction (S3D.h:61-82) */
static void split_axis(float c, float hi, int last, int* i0, float* f) {
	float cc = clampf(c, 0.0f, hi);
	int i = (int)floorf(cc);
	if (i == last) i--;
	*i0 = i;
	*f = cc - (float)i;
}

static float tri(const float* g, int base, int sx, int sy, int sz, float fx, float fy, float fz) {
	/* x lerps, then y, then z (S3D.h:87-97) */
	float a00 = lerpf(g[base], g[base + sx], fx);
	float a10 = lerpf(g[base + sy], g[base + sy + sx], fx);
	float a01 = lerpf(g[base + sz], g[base + sz + sx], fx);
	float a11 = lerpf(g[base + sz + sy], g[base + sz + sy + sx], fx);
	float b0 = lerpf(a00, a10, fy);
	float b1 = lerpf(a01, a11, fy);
	return lerpf(b0, b1, fz);
}

/* S3D.h:55-123.  (i,j,k) are index-space coordinates. */
static void interp(const fs3o* s, const float* U, const float* V, const float* W,
                   float i, float j, float k, float out[3]) {
	int nx = s->nx, ny = s->ny, nz = s->nz;
	int iI, iJ, iK, eI, eJ, eK;
	float fI, fJ, fK, gI, gJ, gK;
	split_axis(i, (float)nx - 1.0f, nx - 1, &iI, &fI);
	split_axis(j, (float)ny - 1.0f, ny - 1, &iJ, &fJ);
	split_axis(k, (float)nz - 1.0f, nz - 1, &iK, &fK);
	split_axis(i + 0.5f, (float)nx, nx, &eI, &gI);
	split_axis(j + 0.5f, (float)ny, ny, &eJ, &gJ);
	split_axis(k + 0.5f, (float)nz, nz, &eK, &gK);
	out[0] = tri(U, IU(s, eI, iJ, iK), 1, nx + 1, (nx + 1) * ny, gI, fJ, fK);
	out[1] = tri(V, IV(s, iI, eJ, iK), 1, nx, nx * (ny + 1), fI, gJ, fK);
	out[2] = tri(W, IW(s, iI, iJ, eK), 1, nx, nx * ny, fI, fJ, gK);
}

void fs3o_interpolate(const fs3o* s, float i, float j, float k, float* out3) {
	interp(s, s->u, s->v, s->w, i, j, k, out3);
}

/* S3D.cpp:41-99 */
void fs3o_reset(fs3o* s) {
	lcg_t g = {1};
	size_t q;
	for (q = 0; q < n_u(s); q++) s->u[q] = lcg_draw(&g);   /* loops are z,y,x with x fastest == linear order */
	for (q = 0; q < n_v(s); q++) s->v[q] = lcg_draw(&g);
	for (q = 0; q < n_w(s); q++) s->w[q] = lcg_draw(&g);
	long cap = 8L * (s->nz - 2) * (s->ny - 2) * (s->nx - 1 - s->nx / 2);
	if (cap < 0) cap = 0;
	free(s->part);
	s->part = (float*)malloc(sizeof(float) * 6 * (size_t)(cap > 0 ? cap : 1));
	long n = 0;
	float cpm = s->cpm;
	for (int z = 1; z < s->nz - 1; z++)
		for (int y = 1; y < s->ny - 1; y++)
			for (int x = s->nx / 2; x < s->nx - 1; x++) {
				float rx = ((float)x - 0.25f) / cpm, ry = ((float)y - 0.25f) / cpm, rz = ((float)z - 0.25f) / cpm;
				float d = 0.5f / cpm;
				for (int a = 0; a <= 1; a++)
					for (int b = 0; b <= 1; b++)
						for (int c = 0; c <= 1; c++) {
							float* p = s->part + 6 * n++;
							p[0] = rx + (float)a * d + lcg_draw(&g) / cpm;
							p[1] = ry + (float)b * d + lcg_draw(&g) / cpm;
							p[2] = rz + (float)c * d + lcg_draw(&g) / cpm;
							interp(s, s->u, s->v, s->w, (float)s->nx * p[0], (float)s->ny * p[1], (float)s->nz * p[2], p + 3);
						}
			}
	s->np = n;
}

/* S3D.cpp:190-232 : Ralston RK3 through the grid, then clamp into the box */
void fs3o_stage_advect(fs3o* s, float dt) {
	float fx = (float)s->nx, fy = (float)s->ny, fz = (float)s->nz;
	const float eps = 0.1f;
	float lox = (-0.5f + eps) / fx, hix = 1.0f + (-0.5f - eps) / fx;
	float loy = (-0.5f + eps) / fy, hiy = 1.0f + (-0.5f - eps) / fy;
	float loz = (-0.5f + eps) / fz, hiz = 1.0f + (-0.5f - eps) / fz;
	for (long i = 0; i < s->np; i++) {
		float* p = s->part + 6 * i;
		float k1[3], k2[3], k3[3];
		interp(s, s->u, s->v, s->w, fx * p[0], fy * p[1], fz * p[2], k1);
		interp(s, s->u, s->v, s->w, fx * (p[0] + 0.5f * dt * k1[0]), fy * (p[1] + 0.5f * dt * k1[1]),
		       fz * (p[2] + 0.5f * dt * k1[2]), k2);
		interp(s, s->u, s->v, s->w, fx * (p[0] + 0.75f * dt * k2[0]), fy * (p[1] + 0.75f * dt * k2[1]),
		       fz * (p[2] + 0.75f * dt * k2[2]), k3);
		float vx = (2.0f / 9.0f) * k1[0] + (3.0f / 9.0f) * k2[0] + (4.0f / 9.0f) * k3[0];
		float vy = (2.0f / 9.0f) * k1[1] + (3.0f / 9.0f) * k2[1] + (4.0f / 9.0f) * k3[1];
		float vz = (2.0f / 9.0f) * k1[2] + (3.0f / 9.0f) * k2[2] + (4.0f / 9.0f) * k3[2];
		p[0] += dt * vx;
		p[1] += dt * vy;
		p[2] += dt * vz;
		p[0] = clampf(p[0], lox, hix);
		p[1] = clampf(p[1], loy, hiy);
		p[2] = clampf(p[2], loz, hiz);
	}
}

/* ---- level set ------------------------------------------------------------------ */

/* S3D.cpp:234-236 */
static float pt_distance(const fs3o* s, float x0, float y0, float z0, float x1, float y1, float z1) {
	return sqrtf((x0 - x1) * (x0 - x1) + (y0 - y1) * (y0 - y1) + (z0 - z1) * (z0 - z1)) - s->radius;
}

/* S3D.cpp:238-253 : offer the upwind neighbour's closest particle to cell (x,y,z).
 * Particle index 0 is never propagated (`otherPt > 0`); strict `<` keeps the incumbent on ties. */
static void offer(fs3o* s, int nb, int x, int y, int z) {
	int cand = s->closest[nb];
	if (cand <= 0) return;
	const float* q = s->part + 6 * (size_t)cand;
	float d = pt_distance(s, (float)x, (float)y, (float)z, q[0] * (float)s->nx, q[1] * (float)s->ny, q[2] * (float)s->nz);
	int c = IC(s, x, y, z);
	if (s->closest[c] < 0 || d < s->phi[c]) {
		s->closest[c] = cand;
		s->phi[c] = d;
	}
}

/* per-particle cell, S3D.cpp:283-291 : (int)roundf(n*X); -1 when outside the grid */
static int particle_cell(const fs3o* s, const float* p) {
	int cx = (int)roundf(p[0] * (float)s->nx);
	int cy = (int)roundf(p[1] * (float)s->ny);
	int cz = (int)roundf(p[2] * (float)s->nz);
	if (cx < 0 || cx >= s->nx || cy < 0 || cy >= s->ny || cz < 0 || cz >= s->nz) return -1;
	return IC(s, cx, cy, cz);
}

/* S3D.cpp:255-420 */
void fs3o_stage_levelset(fs3o* s) {
	size_t nc = n_c(s);
	for (size_t q = 0; q < nc; q++) s->closest[q] = -1;           /* :272-278; phi is NOT cleared */
	for (long i = 0; i < s->np; i++) {                             /* seed, :281-300 */
		const float* p = s->part + 6 * i;
		int c = particle_cell(s, p);
		if (c < 0) continue;
		int cx = c % s->nx, cy = (c / s->nx) % s->ny, cz = c / (s->nx * s->ny);
		float k = pt_distance(s, p[0] * (float)s->nx, p[1] * (float)s->ny, p[2] * (float)s->nz, (float)cx, (float)cy, (float)cz);
		if (s->closest[c] < 0 || s->phi[c] > k) {
			s->closest[c] = (int)i;
			s->phi[c] = k;
		}
	}
	/* 8 Gauss-Seidel sweeps, :306-416.  Sweep m runs x descending when bit0 of m is set,
	 * y descending for bit1, z descending for bit2; x innermost, z outermost; per cell the
	 * upwind x-, y-, z-neighbours are offered in that order. */
	for (int m = 0; m < 8; m++) {
		int dx = (m & 1) ? -1 : 1, dy = (m & 2) ? -1 : 1, dz = (m & 4) ? -1 : 1;
		for (int zz = 0; zz < s->nz; zz++) {
			int z = dz > 0 ? zz : s->nz - 1 - zz;
			for (int yy = 0; yy < s->ny; yy++) {
				int y = dy > 0 ? yy : s->ny - 1 - yy;
				for (int xx = 0; xx < s->nx; xx++) {
					int x = dx > 0 ? xx : s->nx - 1 - xx;
					if (xx != 0) offer(s, IC(s, x - dx, y, z), x, y, z);
					if (yy != 0) offer(s, IC(s, x, y - dy, z), x, y, z);
					if (zz != 0) offer(s, IC(s, x, y, z - dz), x, y, z);
				}
			}
		}
	}
}

void fs3o_get_cell_indices(const fs3o* s, int* cell) {
	for (long i = 0; i < s->np; i++) cell[i] = particle_cell(s, s->part + 6 * i);
}
void fs3o_get_cell_counts(const fs3o* s, int* counts) {
	memset(counts, 0, sizeof(int) * n_c(s));
	for (long i = 0; i < s->np; i++) {
		int c = particle_cell(s, s->part + 6 * i);
		if (c >= 0) counts[c]++;
	}
}

/* ---- extrapolation, S3D.cpp:614-778 ------------------------------------------------ */
void fs3o_extrapolate(float* src, const unsigned char* valid, int xs, int ys, int zs) {
	size_t n = (size_t)xs * ys * zs;
	int* cd = (int*)malloc(sizeof(int) * n);
	const int inf = 1000000000;
	for (size_t q = 0; q < n; q++) cd[q] = valid[q] ? 0 : inf;
	int sx = 1, sy = xs, sz = xs * ys;
	/* six directional min-plus scans: exact L1 distance to the nearest valid node (:641-699) */
	for (int z = 0; z < zs; z++)
		for (int y = 0; y < ys; y++) {
			int* r = cd + (size_t)sy * y + (size_t)sz * z;
			for (int x = 1; x < xs; x++) if (r[x - 1] + 1 < r[x]) r[x] = r[x - 1] + 1;
			for (int x = xs - 2; x >= 0; x--) if (r[x + 1] + 1 < r[x]) r[x] = r[x + 1] + 1;
		}
	for (int z = 0; z < zs; z++) {
		for (int y = 1; y < ys; y++)
			for (int x = 0; x < xs; x++) { int q = x + sy * y + sz * z; if (cd[q - sy] + 1 < cd[q]) cd[q] = cd[q - sy] + 1; }
		for (int y = ys - 2; y >= 0; y--)
			for (int x = 0; x < xs; x++) { int q = x + sy * y + sz * z; if (cd[q + sy] + 1 < cd[q]) cd[q] = cd[q + sy] + 1; }
	}
	for (int z = 1; z < zs; z++)
		for (int q = sz * z; q < sz * (z + 1); q++) if (cd[q - sz] + 1 < cd[q]) cd[q] = cd[q - sz] + 1;
	for (int z = zs - 2; z >= 0; z--)
		for (int q = sz * z; q < sz * (z + 1); q++) if (cd[q + sz] + 1 < cd[q]) cd[q] = cd[q + sz] + 1;
	/* counting sort of nodes by distance, linear (z,y,x) order inside a bucket (:701-740) */
	int maxd = 0;
	for (size_t q = 0; q < n; q++) if (cd[q] > maxd) maxd = cd[q];
	int nb = maxd + 1;
	int* start = (int*)calloc((size_t)nb + 1, sizeof(int));
	for (size_t q = 0; q < n; q++) start[cd[q] + 1]++;
	for (int d = 0; d < nb; d++) start[d + 1] += start[d];
	int* order = (int*)malloc(sizeof(int) * n);
	for (size_t q = 0; q < n; q++) order[start[cd[q]]++] = (int)q;
	/* inside-out fill: mean of the 6-neighbours that are strictly closer, order +x,-x,+y,-y,+z,-z (:742-769) */
	for (size_t t = 0; t < n; t++) {
		int q = order[t];
		int my = cd[q];
		if (my == 0) continue;
		int x = q % xs, y = (q / xs) % ys, z = q / (xs * ys);
		float cnt = 0.0f, sum = 0.0f;
		if (x + 1 < xs && cd[q + sx] < my) { cnt += 1.0f; sum += src[q + sx]; }
		if (x - 1 >= 0 && cd[q - sx] < my) { cnt += 1.0f; sum += src[q - sx]; }
		if (y + 1 < ys && cd[q + sy] < my) { cnt += 1.0f; sum += src[q + sy]; }
		if (y - 1 >= 0 && cd[q - sy] < my) { cnt += 1.0f; sum += src[q - sy]; }
		if (z + 1 < zs && cd[q + sz] < my) { cnt += 1.0f; sum += src[q + sz]; }
		if (z - 1 >= 0 && cd[q - sz] < my) { cnt += 1.0f; sum += src[q - sz]; }
		src[q] = sum / cnt;
	}
	free(cd); free(start); free(order);
}

/* ---- particle-to-grid, S3D.cpp:422-612 --------------------------------------------- */

/* S3D.cpp:1140-1162 */
static void zero_wall_faces(fs3o* s) {
	for (int z = 0; z < s->nz; z++)
		for (int y = 0; y < s->ny; y++) { s->u[IU(s, 0, y, z)] = 0; s->u[IU(s, s->nx, y, z)] = 0; }
	for (int z = 0; z < s->nz; z++)
		for (int x = 0; x < s->nx; x++) { s->v[IV(s, x, 0, z)] = 0; s->v[IV(s, x, s->ny, z)] = 0; }
	for (int y = 0; y < s->ny; y++)
		for (int x = 0; x < s->nx; x++) { s->w[IW(s, x, y, 0)] = 0; s->w[IW(s, x, y, s->nz)] = 0; }
}

/* One face grid of the trilinear-hat splat (S3D.cpp:458-536).  `ax` is the staggered axis:
 * its base index is floorf(p+0.5f), upper guard `<= n[ax]`, no lower guard; the other axes
 * use floorf(p) and 0 <= idx < n.  dims[] are the grid's own extents.  NB the reference
 * indexes uAmts with mZ where mY is meant (S3D.cpp:476) -- identical for cubes, and cubes
 * are all this oracle accepts (fs3o_create). */
static void splat(const fs3o* s, float* grid, float* amt, const int dims[3], int ax, const float pc[3], float vel) {
	int n[3] = {s->nx, s->ny, s->nz};
	int base[3];
	float al[3];
	for (int a = 0; a < 3; a++) {
		float q = (a == ax) ? pc[a] + 0.5f : pc[a];
		base[a] = (int)floorf(q);
		al[a] = q - (float)base[a];
	}
	for (int z = 0; z <= 1; z++)
		for (int y = 0; y <= 1; y++)
			for (int x = 0; x <= 1; x++) {
				int id[3] = {base[0] + x, base[1] + y, base[2] + z};
				int ok = 1;
				for (int a = 0; a < 3; a++) {
					if (a == ax) ok = ok && (id[a] <= n[a]);
					else ok = ok && (0 <= id[a] && id[a] < n[a]);
				}
				if (!ok) continue;
				float wgt = (x > 0 ? al[0] : 1.f - al[0]) * (y > 0 ? al[1] : 1.f - al[1]) * (z > 0 ? al[2] : 1.f - al[2]);
				size_t q = (size_t)id[0] + (size_t)dims[0] * ((size_t)id[1] + (size_t)dims[1] * id[2]);
				grid[q] += wgt * vel;
				amt[q] += wgt;
			}
}

void fs3o_stage_p2g(fs3o* s) {
	size_t nu = n_u(s), nv = n_v(s), nw = n_w(s);
	float* au = (float*)calloc(nu, 4);
	float* av = (float*)calloc(nv, 4);
	float* aw = (float*)calloc(nw, 4);
	memset(s->u, 0, nu * 4); memset(s->v, 0, nv * 4); memset(s->w, 0, nw * 4);
	int du[3] = {s->nx + 1, s->ny, s->nz}, dv[3] = {s->nx, s->ny + 1, s->nz}, dw[3] = {s->nx, s->ny, s->nz + 1};
	for (long i = 0; i < s->np; i++) {
		const float* p = s->part + 6 * i;
		float pc[3] = {p[0] * s->cpm, p[1] * s->cpm, p[2] * s->cpm};   /* CellsPerMeter on every axis (:442-444) */
		if ((double)pc[0] < -0.5 || (double)pc[0] > (s->nx + 0.5) || (double)pc[1] < -0.5 || (double)pc[1] > (s->ny + 0.5) ||
		    (double)pc[2] < -0.5 || (double)pc[2] > (s->nz + 0.5))
			continue;
		splat(s, s->u, au, du, 0, pc, p[3]);
		splat(s, s->v, av, dv, 1, pc, p[4]);
		splat(s, s->w, aw, dw, 2, pc, p[5]);
	}
	/* normalise, :540-561 */
	const float tiny = 1.401298464324817e-45f; /* numeric_limits<float>::denorm_min() */
	for (size_t q = 0; q < nu; q++) s->u[q] = s->u[q] / (tiny + au[q]);
	for (size_t q = 0; q < nv; q++) s->v[q] = s->v[q] / (tiny + av[q]);
	for (size_t q = 0; q < nw; q++) s->w[q] = s->w[q] / (tiny + aw[q]);
	/* validity, :565-599 */
	unsigned char* ku = (unsigned char*)malloc(nu);
	unsigned char* kv = (unsigned char*)malloc(nv);
	unsigned char* kw = (unsigned char*)malloc(nw);
	for (size_t q = 0; q < nu; q++) ku[q] = au[q] > 0.01f;
	for (size_t q = 0; q < nv; q++) kv[q] = av[q] > 0.01f;
	for (size_t q = 0; q < nw; q++) kw[q] = aw[q] > 0.01f;
	zero_wall_faces(s);
	for (int z = 0; z < s->nz; z++)
		for (int y = 0; y < s->ny; y++) { ku[IU(s, 0, y, z)] = 1; ku[IU(s, s->nx, y, z)] = 1; }
	for (int z = 0; z < s->nz; z++)
		for (int x = 0; x < s->nx; x++) { kv[IV(s, x, 0, z)] = 1; kv[IV(s, x, s->ny, z)] = 1; }
	for (int y = 0; y < s->ny; y++)
		for (int x = 0; x < s->nx; x++) { kw[IW(s, x, y, 0)] = 1; kw[IW(s, x, y, s->nz)] = 1; }
	fs3o_extrapolate(s->u, ku, s->nx + 1, s->ny, s->nz);
	fs3o_extrapolate(s->v, kv, s->nx, s->ny + 1, s->nz);
	fs3o_extrapolate(s->w, kw, s->nx, s->ny, s->nz + 1);
	free(au); free(av); free(aw); free(ku); free(kv); free(kw);
}

/* S3D.cpp:780-788 : every V entry, wall faces included */
void fs3o_stage_bodyforces(fs3o* s, float dt) {
	float g = s->gy * dt;
	size_t nv = n_v(s);
	for (size_t q = 0; q < nv; q++) s->v[q] += g;
}

/* ---- pressure projection, S3D.cpp:790-1093 ------------------------------------------- */

static void record_residual(fs3o* s, const double* p) {
	/* r = b - (diag p_c - sum of fluid-neighbour p) over phi<0 cells; not in the reference (it never
	 * measures a residual) -- instrumentation that leaves p untouched. */
	int nx = s->nx, ny = s->ny, nz = s->nz, sy = nx, sz = nx * ny;
	double l2 = 0, li = 0;
	for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {
		int c = IC(s, x, y, z);
		if (s->phi[c] >= 0.0f) continue;
		double nms = 0;
		if (x != 0 && s->phi[c - 1] < 0.0f) nms -= p[c - 1];
		if (x != nx - 1 && s->phi[c + 1] < 0.0f) nms -= p[c + 1];
		if (y != 0 && s->phi[c - sy] < 0.0f) nms -= p[c - sy];
		if (y != ny - 1 && s->phi[c + sy] < 0.0f) nms -= p[c + sy];
		if (z != 0 && s->phi[c - sz] < 0.0f) nms -= p[c - sz];
		if (z != nz - 1 && s->phi[c + sz] < 0.0f) nms -= p[c + sz];
		double r = s->b[c] - (s->diag[c] * p[c] + nms);
		l2 += r * r;
		if (fabs(r) > li) li = fabs(r);
	}
	s->res_l2[s->n_res] = sqrt(l2);
	s->res_linf[s->n_res] = li;
	s->n_res++;
}

#define DEFINE_PROJECT(NAME, REAL)                                                                          \
static void NAME(fs3o* s, float dt) {                                                                       \
	int nx = s->nx, ny = s->ny, nz = s->nz, sy = nx, sz = nx * ny;                                          \
	size_t nc = n_c(s);                                                                                     \
	const REAL max_ratio = (REAL)1000.0;                       /* :823 */                                   \
	REAL* b = (REAL*)malloc(sizeof(REAL) * nc);                                                             \
	REAL* p = (REAL*)calloc(nc, sizeof(REAL));                                                              \
	REAL* dg = (REAL*)calloc(nc, sizeof(REAL));                                                             \
	REAL dx = (REAL)1.0 / s->cpm;                              /* :834 */                                   \
	REAL scale = -dx * s->rho / dt;                            /* :835 */                                   \
	const float *U = s->u, *V = s->v, *W = s->w, *phi = s->phi;                                            \
	/* I. right-hand side; wall faces count as velocity 0 (:837-849) */                                     \
	for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {                \
		REAL xp = (x == nx - 1) ? (REAL)0 : (REAL)U[IU(s, x + 1, y, z)];                                    \
		REAL xm = (x == 0) ? (REAL)0 : (REAL)U[IU(s, x, y, z)];                                             \
		REAL yp = (y == ny - 1) ? (REAL)0 : (REAL)V[IV(s, x, y + 1, z)];                                    \
		REAL ym = (y == 0) ? (REAL)0 : (REAL)V[IV(s, x, y, z)];                                             \
		REAL zp = (z == nz - 1) ? (REAL)0 : (REAL)W[IW(s, x, y, z + 1)];                                    \
		REAL zm = (z == 0) ? (REAL)0 : (REAL)W[IW(s, x, y, z)];                                             \
		b[IC(s, x, y, z)] = scale * (xp - xm + yp - ym + zp - zm);                                          \
	}                                                                                                       \
	/* diagonal: in-domain neighbours + ghost-fluid term for air neighbours (:861-917) */                   \
	for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++) for (int x = 0; x < nx; x++) {                \
		int c = IC(s, x, y, z);                                                                             \
		if (phi[c] >= 0.0f) continue;                                                                       \
		REAL d = 0;                                                                                         \
		const int has[6] = {x != 0, x != nx - 1, y != 0, y != ny - 1, z != 0, z != nz - 1};                 \
		const int off[6] = {-1, 1, -sy, sy, -sz, sz};                                                       \
		for (int k = 0; k < 6; k++) {                                                                       \
			if (!has[k]) continue;                                                                          \
			d += 1;                                                                                         \
			float pn = phi[c + off[k]];                                                                     \
			if (pn > 0.0f) {                                                                                \
				REAL r = -(REAL)pn / (REAL)phi[c];                                                          \
				d += r < 0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);                                     \
			}                                                                                               \
		}                                                                                                   \
		dg[c] = d;                                                                                          \
	}                                                                                                       \
	if (sizeof(REAL) == 8) { memcpy(s->b, b, nc * 8); memcpy(s->diag, dg, nc * 8); }                        \
	else for (size_t q = 0; q < nc; q++) { s->b[q] = (double)b[q]; s->diag[q] = (double)dg[q]; }            \
	s->n_res = 0;                                                                                           \
	/* II. red-black SOR, colour 0 first (:944-1001) */                                                     \
	REAL omega = (REAL)2 - (REAL)3.16343 / nx;                                                              \
	for (int it = 0; it < s->sor_iters; it++) {                                                             \
		for (int colour = 0; colour <= 1; colour++)                                                         \
			for (int z = 0; z < nz; z++) for (int y = 0; y < ny; y++)                                       \
				for (int x = (colour + y + z) & 1; x < nx; x += 2) {                                        \
					int c = IC(s, x, y, z);                                                                 \
					if (phi[c] >= 0.0f) continue;                                                           \
					REAL nms = 0;                                                                           \
					if (x != 0 && phi[c - 1] < 0.0f) nms -= p[c - 1];                                       \
					if (x != nx - 1 && phi[c + 1] < 0.0f) nms -= p[c + 1];                                  \
					if (y != 0 && phi[c - sy] < 0.0f) nms -= p[c - sy];                                     \
					if (y != ny - 1 && phi[c + sy] < 0.0f) nms -= p[c + sy];                                \
					if (z != 0 && phi[c - sz] < 0.0f) nms -= p[c - sz];                                     \
					if (z != nz - 1 && phi[c + sz] < 0.0f) nms -= p[c + sz];                                \
					p[c] = ((REAL)1 - omega) * p[c] + omega * (b[c] - nms) / dg[c];                         \
				}                                                                                           \
		if (s->record_res) {                                                                                \
			if (sizeof(REAL) == 8) record_residual(s, (const double*)(const void*)p);                       \
			else { for (size_t q = 0; q < nc; q++) s->p[q] = (double)p[q]; record_residual(s, s->p); }      \
		}                                                                                                   \
	}                                                                                                       \
	/* III. subtract the pressure gradient, four ghost-fluid cases per face (:1012-1084) */                 \
	zero_wall_faces(s);                                                                                     \
	scale = dt / (s->rho * dx);                                                                             \
	for (int axis = 0; axis < 3; axis++) {                                                                  \
		float* G = axis == 0 ? s->u : (axis == 1 ? s->v : s->w);                                            \
		int lx = nx - (axis == 0), ly = ny - (axis == 1), lz = nz - (axis == 2);                            \
		int step = axis == 0 ? 1 : (axis == 1 ? sy : sz);                                                   \
		for (int z = 0; z < lz; z++) for (int y = 0; y < ly; y++) for (int x = 0; x < lx; x++) {            \
			int c = IC(s, x, y, z);                                                                         \
			int f = axis == 0 ? IU(s, x + 1, y, z) : (axis == 1 ? IV(s, x, y + 1, z) : IW(s, x, y, z + 1)); \
			REAL lo = phi[c], hi = phi[c + step];                                                           \
			if (lo < 0 && hi < 0) G[f] = (float)(G[f] - scale * (p[c + step] - p[c]));                      \
			else if (lo < 0 && hi >= 0) {                                                                   \
				REAL r = -hi / lo; r = r < 0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);                   \
				G[f] = (float)(G[f] + scale * (1 + r) * p[c]);                                              \
			} else if (lo >= 0 && hi < 0) {                                                                 \
				REAL r = -lo / hi; r = r < 0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);                   \
				G[f] = (float)(G[f] - scale * (1 + r) * p[c + step]);                                       \
			} else G[f] = 0;                                                                                \
		}                                                                                                   \
	}                                                                                                       \
	for (size_t q = 0; q < nc; q++) s->p[q] = (double)p[q];                                                 \
	free(b); free(p); free(dg);                                                                             \
}

DEFINE_PROJECT(project_f64, double)
DEFINE_PROJECT(project_f32, float)

void fs3o_stage_project(fs3o* s, float dt) { project_f64(s, dt); }
void fs3o_stage_project_f32(fs3o* s, float dt) { project_f32(s, dt); }

/* ---- PIC/FLIP blend, S3D.cpp:112-180 -------------------------------------------------- */

static float flip_alpha(const fs3o* s, float dt) {
	return clampf(6 * dt * s->nu * s->cpm * s->cpm, 0.0f, 1.0f);   /* :114 */
}

void fs3o_stage_g2p(fs3o* s, float dt, const float* old_u, const float* old_v, const float* old_w) {
	float alpha = flip_alpha(s, dt);
	size_t nu = n_u(s), nv = n_v(s), nw = n_w(s);
	float* du = (float*)malloc(nu * 4);
	float* dv = (float*)malloc(nv * 4);
	float* dw = (float*)malloc(nw * 4);
	for (size_t q = 0; q < nu; q++) du[q] = s->u[q] - (1.0f - alpha) * old_u[q];   /* :148-156 */
	for (size_t q = 0; q < nv; q++) dv[q] = s->v[q] - (1.0f - alpha) * old_v[q];
	for (size_t q = 0; q < nw; q++) dw[q] = s->w[q] - (1.0f - alpha) * old_w[q];
	for (long i = 0; i < s->np; i++) {                                            /* :163-176 */
		float* p = s->part + 6 * i;
		float d[3];
		interp(s, du, dv, dw, (float)s->nx * p[0], (float)s->ny * p[1], (float)s->nz * p[2], d);
		p[3] = (1.0f - alpha) * p[3] + d[0];
		p[4] = (1.0f - alpha) * p[4] + d[1];
		p[5] = (1.0f - alpha) * p[5] + d[2];
		if (p[3] > 100000) p[3] = 1000000;                                        /* :172-175 */
	}
	free(du); free(dv); free(dw);
}

/* S3D.cpp:101-188 */
void fs3o_step(fs3o* s, float dt) {
	dt = clampf(dt, 0.0f, 1.0f / 15.0f);
	dt = 0.01f;                                                                   /* :104 */
	fs3o_stage_advect(s, dt);
	fs3o_stage_levelset(s);
	fs3o_stage_p2g(s);
	size_t nu = n_u(s), nv = n_v(s), nw = n_w(s);
	float* ou = (float*)malloc(nu * 4);
	float* ov = (float*)malloc(nv * 4);
	float* ow = (float*)malloc(nw * 4);
	memcpy(ou, s->u, nu * 4); memcpy(ov, s->v, nv * 4); memcpy(ow, s->w, nw * 4);
	fs3o_stage_bodyforces(s, dt);
	fs3o_stage_project(s, dt);
	fs3o_stage_g2p(s, dt, ou, ov, ow);
	free(ou); free(ov); free(ow);
}

/* ---- lifecycle / accessors ------------------------------------------------------------- */

fs3o* fs3o_create(int nx, int ny, int nz, float cpm) {
	if (nx < 4 || nx != ny || ny != nz || cpm != (float)nx) return NULL;  /* cubes with CellsPerMeter == n only (see splat) */
	fs3o* s = (fs3o*)calloc(1, sizeof(fs3o));
	s->nx = nx; s->ny = ny; s->nz = nz; s->cpm = cpm;
	s->rho = 1000.0f; s->gy = -9.81f; s->nu = 8.90f * 1e-4f / 1000.0f; s->radius = 1.0f;  /* S3D.h:135-145 */
	s->u = (float*)malloc(n_u(s) * 4); s->v = (float*)malloc(n_v(s) * 4); s->w = (float*)malloc(n_w(s) * 4);
	s->phi = (float*)calloc(n_c(s), 4);   /* the reference leaves m_Phi uninitialised; every cell is written by the first level set */
	s->closest = (int*)malloc(n_c(s) * sizeof(int));
	s->p = (double*)calloc(n_c(s), 8); s->b = (double*)calloc(n_c(s), 8); s->diag = (double*)calloc(n_c(s), 8);
	s->sor_iters = 100;
	s->res_l2 = (double*)calloc(100000, 8); s->res_linf = (double*)calloc(100000, 8);
	fs3o_reset(s);
	return s;
}
void fs3o_destroy(fs3o* s) {
	if (!s) return;
	free(s->u); free(s->v); free(s->w); free(s->phi); free(s->part); free(s->closest);
	free(s->p); free(s->b); free(s->diag); free(s->res_l2); free(s->res_linf); free(s);
}
long fs3o_num_particles(const fs3o* s) { return s->np; }
void fs3o_get_particles(const fs3o* s, float* out6) { memcpy(out6, s->part, (size_t)s->np * 24); }
void fs3o_set_particles(fs3o* s, const float* in6, long count) {
	free(s->part);
	s->part = (float*)malloc((size_t)(count > 0 ? count : 1) * 24);
	memcpy(s->part, in6, (size_t)count * 24);
	s->np = count;
}
void fs3o_get_velocity(const fs3o* s, float* u, float* v, float* w) {
	if (u) memcpy(u, s->u, n_u(s) * 4);
	if (v) memcpy(v, s->v, n_v(s) * 4);
	if (w) memcpy(w, s->w, n_w(s) * 4);
}
void fs3o_set_velocity(fs3o* s, const float* u, const float* v, const float* w) {
	if (u) memcpy(s->u, u, n_u(s) * 4);
	if (v) memcpy(s->v, v, n_v(s) * 4);
	if (w) memcpy(s->w, w, n_w(s) * 4);
}
void fs3o_get_phi(const fs3o* s, float* phi) { memcpy(phi, s->phi, n_c(s) * 4); }
void fs3o_set_phi(fs3o* s, const float* phi) { memcpy(s->phi, phi, n_c(s) * 4); }
void fs3o_get_closest(const fs3o* s, int* c) { memcpy(c, s->closest, n_c(s) * sizeof(int)); }
void fs3o_get_pressure(const fs3o* s, double* p) { memcpy(p, s->p, n_c(s) * 8); }
void fs3o_get_rhs_diag(const fs3o* s, double* b, double* diag) {
	if (b) memcpy(b, s->b, n_c(s) * 8);
	if (diag) memcpy(diag, s->diag, n_c(s) * 8);
}
int fs3o_get_residuals(const fs3o* s, double* l2, double* linf, int cap) {
	int n = s->n_res < cap ? s->n_res : cap;
	if (l2) memcpy(l2, s->res_l2, (size_t)n * 8);
	if (linf) memcpy(linf, s->res_linf, (size_t)n * 8);
	return s->n_res;
}
void fs3o_set_sor_iterations(fs3o* s, int iters) { s->sor_iters = iters < 0 ? 0 : (iters > 100000 ? 100000 : iters); }
void fs3o_set_record_residuals(fs3o* s, int on) { s->record_res = on; }

/* S3D.cpp:1095-1138 : divergence (no dx scaling) over phi<0 cells */
void fs3o_divergence(const fs3o* s, double* maxdiv, double* l2) {
	double m = 0, acc = 0;
	for (int z = 0; z < s->nz; z++) for (int y = 0; y < s->ny; y++) for (int x = 0; x < s->nx; x++) {
		if (s->phi[IC(s, x, y, z)] >= 0.0f) continue;
		float d = s->v[IV(s, x, y + 1, z)] - s->v[IV(s, x, y, z)] + s->u[IU(s, x + 1, y, z)] - s->u[IU(s, x, y, z)] +
		          s->w[IW(s, x, y, z + 1)] - s->w[IW(s, x, y, z)];
		if (d > m) m = d;
		acc += (double)(d * d);
	}
	*maxdiv = m; *l2 = sqrt(acc);
}
