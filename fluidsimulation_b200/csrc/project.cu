// Project (S3D.cpp:790-1093) on sm_100a: right-hand side + ghost-fluid diagonal, red-black SOR,
// pressure-gradient update.  The pressure system is stored COLOUR-SPLIT (see RBLayout): a half sweep
// streams one colour's p/b/diag contiguously with 128-bit accesses and reads the other colour as a
// 7-point stencil whose y/z neighbours sit at the same index in the adjacent row/plane.  Neighbour
// "is fluid" tests are folded into one mask byte per cell so phi is never re-read in the 200 sweeps.
// REAL = double reproduces the reference's fp64 solve bit for bit (each half sweep is order
// independent; same IEEE operations, no FMA contraction); REAL = float is the fast mode.
#include "fs3_internal.cuh"
#include <string>
#include <algorithm>
#include <cstdlib>

template <typename T> struct Ops;
template <> struct Ops<double> {
	static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
	static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
	static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
	static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
};
template <> struct Ops<float> {
	static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
	static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
	static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
	static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
};

template <typename REAL> struct Sys {
	REAL *p[2], *b[2], *dg[2];
	unsigned char* mask[2];
	RBLayout rb;
};

// ---- I. rhs, diagonal, masks (S3D.cpp:833-917); also clears p --------------------------------------
template <typename REAL>
__global__ void __launch_bounds__(256) k_project_setup(Sys<REAL> sy, const float* __restrict__ U, const float* __restrict__ V,
                                                       const float* __restrict__ W, const float* __restrict__ phi, REAL scale, float gdt,
                                                       int add_g, REAL max_ratio, size_t c0, size_t c1) {
	typedef Ops<REAL> O;
	const int nx = sy.rb.nx, ny = sy.rb.ny, nz = sy.rb.nz;
	size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= c1) return;
	int x = (int)(c % nx), y = (int)((c / nx) % ny), z = (int)(c / ((size_t)nx * ny));
	size_t iu = (size_t)x + (size_t)(nx + 1) * ((size_t)y + (size_t)ny * z);
	size_t iv = (size_t)x + (size_t)nx * ((size_t)y + (size_t)(ny + 1) * z);
	size_t iw = c;
	// wall faces count as solid velocity 0 (S3D.cpp:840-845)
	REAL xp = x == nx - 1 ? (REAL)0 : (REAL)U[iu + 1];
	REAL xm = x == 0 ? (REAL)0 : (REAL)U[iu];
	float vhi = V[iv + nx], vlo = V[iv];
	if (add_g) { vhi = __fadd_rn(vhi, gdt); vlo = __fadd_rn(vlo, gdt); }   // AddBodyForces folded in (S3D.cpp:783-787)
	REAL yp = y == ny - 1 ? (REAL)0 : (REAL)vhi;
	REAL ym = y == 0 ? (REAL)0 : (REAL)vlo;
	REAL zp = z == nz - 1 ? (REAL)0 : (REAL)W[iw + (size_t)nx * ny];
	REAL zm = z == 0 ? (REAL)0 : (REAL)W[iw];
	REAL div = O::sub(O::add(O::sub(O::add(O::sub(xp, xm), yp), ym), zp), zm);
	REAL bval = O::mul(scale, div);
	float pc = phi[c];
	unsigned m = 0;
	REAL d = 0;
	if (pc < 0.0f) {
		m = 64 | 128;   // bit7: no air neighbour, i.e. diag == number of fluid neighbours exactly (cleared below)
		const int sY = nx, sZ = nx * ny;
		const bool has[6] = {x != 0, x != nx - 1, y != 0, y != ny - 1, z != 0, z != nz - 1};
		const long long off[6] = {-1, 1, -sY, sY, -(long long)sZ, (long long)sZ};
#pragma unroll
		for (int k = 0; k < 6; k++) {
			if (!has[k]) continue;
			d = O::add(d, (REAL)1);
			float pn = phi[(long long)c + off[k]];
			if (pn < 0.0f) m |= 1u << k;
			if (!(pn < 0.0f)) m &= ~128u;
			if (pn > 0.0f) {
				REAL r = O::div(-(REAL)pn, (REAL)pc);
				r = r < (REAL)0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);
				d = O::add(d, r);
			}
		}
	}
	int col = (x + y + z) & 1;
	size_t q = sy.rb.at(x >> 1, y, z);
	sy.b[col][q] = bval;
	sy.dg[col][q] = d;
	sy.p[col][q] = (REAL)0;
	sy.mask[col][q] = (unsigned char)m;
}

// ---- II. one colour of red-black SOR (S3D.cpp:948-1001) --------------------------------------------
template <typename REAL, int VEC> struct VecT;
template <> struct VecT<double, 2> { typedef double2 type; typedef unsigned short mtype; };
template <> struct VecT<float, 4> { typedef float4 type; typedef unsigned int mtype; };
// 128-bit loads / stores into register arrays (component-wise, so that the arrays never get an address and stay in registers)
__device__ __forceinline__ void ldv(const double* p, double (&a)[2]) { const double2 t = *reinterpret_cast<const double2*>(p); a[0] = t.x; a[1] = t.y; }
__device__ __forceinline__ void ldv(const float* p, float (&a)[4]) { const float4 t = *reinterpret_cast<const float4*>(p); a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w; }
// streaming (evict-first) loads for the operands that are read once per half sweep -- b, diag, mask, list -- so that they do
// not push the two colours of p, which every half sweep reads again, out of the 126 MB L2
__device__ __forceinline__ void ldv_stream(const double* p, double (&a)[2]) { const double2 t = __ldcs(reinterpret_cast<const double2*>(p)); a[0] = t.x; a[1] = t.y; }
__device__ __forceinline__ void ldv_stream(const float* p, float (&a)[4]) { const float4 t = __ldcs(reinterpret_cast<const float4*>(p)); a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w; }
// L2 loads: inside ONE launch that runs many half sweeps, p written by other SMs (or by the neighbouring GPU) in the previous half
// sweep must not be served from this SM's L1
__device__ __forceinline__ void ldv_cg(const double* p, double (&a)[2]) { const double2 t = __ldcg(reinterpret_cast<const double2*>(p)); a[0] = t.x; a[1] = t.y; }
__device__ __forceinline__ void ldv_cg(const float* p, float (&a)[4]) { const float4 t = __ldcg(reinterpret_cast<const float4*>(p)); a[0] = t.x; a[1] = t.y; a[2] = t.z; a[3] = t.w; }
// system-scope loads: operands that live in (or were just written by) a neighbouring GPU
__device__ __forceinline__ void ldv_sys(const double* p, double (&a)[2]) { asm volatile("ld.relaxed.sys.global.v2.f64 {%0, %1}, [%2];\n" : "=d"(a[0]), "=d"(a[1]) : "l"(p) : "memory"); }
__device__ __forceinline__ void ldv_sys(const float* p, float (&a)[4]) {
	asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];\n" : "=f"(a[0]), "=f"(a[1]), "=f"(a[2]), "=f"(a[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ double ld_sys(const double* p) { double v; asm volatile("ld.relaxed.sys.global.f64 %0, [%1];\n" : "=d"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ float ld_sys(const float* p) { float v; asm volatile("ld.relaxed.sys.global.f32 %0, [%1];\n" : "=f"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void stv(double* p, const double (&a)[2]) { *reinterpret_cast<double2*>(p) = make_double2(a[0], a[1]); }
__device__ __forceinline__ void stv(float* p, const float (&a)[4]) { *reinterpret_cast<float4*>(p) = make_float4(a[0], a[1], a[2], a[3]); }

// The sweep runs over a compacted list of the colour's ACTIVE vectors (VEC consecutive colour cells of which
// at least one is fluid), built once per projection (200 sweeps reuse it): every resident warp does useful
// work, air is never touched, and consecutive list entries are consecutive in memory wherever the fluid
// is, so accesses stay coalesced.  Entry = vector index | (x parity of the colour in that row) << 31.
// Interior cells (mask bit7) have diag == popcount(fluid neighbours), so dg is only read near the surface.
// Sharded sweep (SHARD): the list holds the vectors of this rank's slab.  The kernel first waits until both z-neighbours
// have published the previous half sweep (their boundary plane of the OTHER colour was stored straight into this rank's
// ghost planes), stores the updated vectors of its own boundary planes into the neighbours' arrays over NVLink while it
// sweeps, and the last block to finish publishes the epoch to both neighbours: compute and halo exchange are one kernel.
struct SorShard {
	void* peer_p[2];          // this colour's array of the lower / upper neighbour (nullptr: no neighbour)
	uint32_t* peer_flag[2];   // where that neighbour expects this rank's epoch
	const uint32_t* my_flag;  // [0] written by the lower, [1] by the upper neighbour
	uint32_t wait_epoch, signal_epoch;
	size_t lo_begin, lo_end, hi_begin, hi_end;   // element ranges of the slab's first / last plane in a colour array
	uint32_t* done;           // block counter
	const uint32_t *n_lo, *n_hi;   // list positions: entries [0, n_lo) lie in the slab's first plane, [n_hi, n) in its last
	int* flags;
	int nospin;
};
// one list entry: VEC consecutive cells of one colour.  PUSH: also store the result into the neighbours' ghost planes.
// The colour-dependent pointers are resolved once per thread with selects: indexing the kernel-parameter structs with a
// run-time colour would make the compiler copy them to local memory.
template <typename REAL> struct SorPtrs {
	REAL* own; const REAL* other; const REAL* b; const REAL* dg; const unsigned char* mask;
	size_t hp, plane;
	REAL *peer_lo, *peer_hi;
	size_t lo_begin, lo_end, hi_begin, hi_end;
};
template <typename REAL>
__device__ __forceinline__ SorPtrs<REAL> sor_ptrs(const Sys<REAL>& sy, int colour, const SorShard& sh) {
	SorPtrs<REAL> sp;
	sp.own = colour ? sy.p[1] : sy.p[0]; sp.other = colour ? sy.p[0] : sy.p[1];
	sp.b = colour ? sy.b[1] : sy.b[0]; sp.dg = colour ? sy.dg[1] : sy.dg[0]; sp.mask = colour ? sy.mask[1] : sy.mask[0];
	sp.hp = (size_t)sy.rb.hp; sp.plane = sy.rb.plane;
	sp.peer_lo = reinterpret_cast<REAL*>(sh.peer_p[0]); sp.peer_hi = reinterpret_cast<REAL*>(sh.peer_p[1]);
	sp.lo_begin = sh.lo_begin; sp.lo_end = sh.lo_end; sp.hi_begin = sh.hi_begin; sp.hi_end = sh.hi_end;
	return sp;
}
// EDGE (two-plane halos, red entries next to a neighbour): the other colour's values of the neighbour's planes are read from the
// NEIGHBOUR's array over NVLink (same layout, so the same offsets): sp.peer_lo / peer_hi are its other-colour arrays, sp.lo_begin /
// hi_begin the first element of this rank's first / last plane.
template <typename REAL, int VEC>
__device__ __forceinline__ typename VecT<REAL, VEC>::mtype sor_mask(const SorPtrs<REAL>& sp, uint32_t ent) {
	return __ldcs(reinterpret_cast<const typename VecT<REAL, VEC>::mtype*>(sp.mask + (size_t)(ent & 0x7fffffffu) * VEC));
}
// (mm: the entry's mask word, loaded by the caller -- ahead of time where the loop can)
template <typename REAL, int VEC, bool PUSH, bool CG = false, bool EDGE = false>
__device__ __forceinline__ void sor_vector_m(const SorPtrs<REAL>& sp, REAL omega, REAL one_minus_omega, uint32_t ent, typename VecT<REAL, VEC>::mtype mm) {
	typedef Ops<REAL> O;
	typedef typename VecT<REAL, VEC>::type V;
	typedef typename VecT<REAL, VEC>::mtype M;
	const REAL* __restrict__ o = sp.other;
	const M nodiag = VEC == 2 ? (M)0x8080 : (M)0x80808080u;
	const int s = (int)(ent >> 31);   // x parity of this colour in this row
	const size_t base = (size_t)(ent & 0x7fffffffu) * VEC;
	REAL own[VEC], bb[VEC], dd[VEC], xc[VEC], ym[VEC], yp[VEC], zm[VEC], zp[VEC];
	REAL edge;
	if (EDGE) {
		const REAL *oc = o, *ozm = o, *ozp = o;   // in-plane, plane below, plane above
		if (sp.peer_lo && base < sp.lo_begin + sp.plane) { ozm = sp.peer_lo; if (base < sp.lo_begin) oc = sp.peer_lo; }
		if (sp.peer_hi && base >= sp.hi_begin) { ozp = sp.peer_hi; if (base >= sp.hi_begin + sp.plane) oc = sp.peer_hi; }
		ldv_cg(sp.own + base, own);
		ldv_sys(oc + base, xc);
		ldv_sys(oc + base - sp.hp, ym); ldv_sys(oc + base + sp.hp, yp);
		ldv_sys(ozm + base - sp.plane, zm); ldv_sys(ozp + base + sp.plane, zp);
		edge = ld_sys(s == 0 ? oc + base - 1 : oc + base + VEC);
	} else if (CG) {
		ldv_cg(sp.own + base, own);
		ldv_cg(o + base, xc);
		ldv_cg(o + base - sp.hp, ym); ldv_cg(o + base + sp.hp, yp);
		ldv_cg(o + base - sp.plane, zm); ldv_cg(o + base + sp.plane, zp);
		edge = __ldcg(s == 0 ? o + base - 1 : o + base + VEC);
	} else {
		ldv(sp.own + base, own);
		ldv(o + base, xc);
		ldv(o + base - sp.hp, ym); ldv(o + base + sp.hp, yp);
		ldv(o + base - sp.plane, zm); ldv(o + base + sp.plane, zp);
		edge = s == 0 ? o[base - 1] : o[base + VEC];
	}
	ldv_stream(sp.b + base, bb);
#pragma unroll
	for (int e = 0; e < VEC; e++) dd[e] = (REAL)1;
	if ((mm & nodiag) != nodiag) ldv_stream(sp.dg + base, dd);
#pragma unroll
	for (int e = 0; e < VEC; e++) {
		const unsigned m = (unsigned)(mm >> (8 * e)) & 0xffu;
		if (!(m & 64u)) continue;
		// other-colour x neighbours: left at index e-1+s, right at e+s
		const REAL L = s ? xc[e] : (e == 0 ? edge : xc[e > 0 ? e - 1 : 0]);
		const REAL R = s ? (e == VEC - 1 ? edge : xc[e < VEC - 1 ? e + 1 : e]) : xc[e];
		REAL nms = 0;                                   // neighborMinusSum, order x-,x+,y-,y+,z-,z+ (S3D.cpp:964-993)
		if (m & 1u) nms = O::sub(nms, L);
		if (m & 2u) nms = O::sub(nms, R);
		if (m & 4u) nms = O::sub(nms, ym[e]);
		if (m & 8u) nms = O::sub(nms, yp[e]);
		if (m & 16u) nms = O::sub(nms, zm[e]);
		if (m & 32u) nms = O::sub(nms, zp[e]);
		const REAL dg = (m & 128u) ? (REAL)__popc(m & 63u) : dd[e];
		// p = (1-omega) p + omega (b - nms) / diag   (S3D.cpp:996)
		own[e] = O::add(O::mul(one_minus_omega, own[e]), O::div(O::mul(omega, O::sub(bb[e], nms)), dg));
	}
	stv(sp.own + base, own);
	if (PUSH) {
		if (sp.peer_lo && base >= sp.lo_begin && base < sp.lo_end) stv(sp.peer_lo + base, own);
		if (sp.peer_hi && base >= sp.hi_begin && base < sp.hi_end) stv(sp.peer_hi + base, own);
	}
}

template <typename REAL, int VEC, bool PUSH, bool CG = false, bool EDGE = false>
__device__ __forceinline__ void sor_vector(const SorPtrs<REAL>& sp, REAL omega, REAL one_minus_omega, uint32_t ent) {
	sor_vector_m<REAL, VEC, PUSH, CG, EDGE>(sp, omega, one_minus_omega, ent, sor_mask<REAL, VEC>(sp, ent));
}

// SHARD: BOUNDARY FIRST.  The list is in memory order, so the active vectors of the slab's first plane are its first n_lo
// entries and those of its last plane the entries from n_hi on (both read from the scan of the activity flags).  The
// lowest-numbered blocks sweep these two planes first, storing every result into the neighbour's ghost plane as well,
// and the last of them to finish raises the neighbours' epoch flags; only then does the grid turn to the interior.  The
// neighbours therefore get this half sweep's halo while the interior is still being swept, and the halo this rank waits
// for at the start of the NEXT half sweep was sent a whole interior sweep earlier: the NVLink latency is off the critical
// path.  (No write-after-read hazard: ghost planes are only read by the boundary vectors, i.e. before the signal that
// allows the neighbour to overwrite them.)
#ifndef FS3_SOR_MINB
#define FS3_SOR_MINB 4
#endif
#ifndef FS3_SOR_THREADS
#define FS3_SOR_THREADS 256
#endif
template <typename REAL, int VEC, bool SHARD>
__global__ void __launch_bounds__(FS3_SOR_THREADS, FS3_SOR_MINB) k_sor_half(Sys<REAL> sy, int colour, REAL omega, REAL one_minus_omega,
                                                  const uint32_t* __restrict__ list, const uint32_t* __restrict__ count, SorShard sh) {
	const uint32_t n = __ldg(count);
	const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x, G = gridDim.x * blockDim.x;
	const SorPtrs<REAL> sp = sor_ptrs<REAL>(sy, colour, sh);
	if (!SHARD) {
		// Programmatic dependent launch: the next half sweep may be scheduled as soon as every CTA of this one has got
		// here (its CTAs take the SM slots this grid frees while it drains and wait below), and this grid waits until the
		// previous half sweep has completed and its stores are visible.  The list, the mask and b are constant during
		// the solve, so the first entry is fetched ahead of the wait.
		asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
		// A thread's entries are serial and each is a chain of dependent loads (list -> mask -> p): the list entry is fetched two
		// entries ahead and the mask one ahead, so that only the p loads remain on the chain.
		typedef typename VecT<REAL, VEC>::mtype M;
		uint32_t e0 = gtid < n ? __ldg(list + gtid) : 0u, e1 = gtid + G < n ? __ldg(list + gtid + G) : 0u;
		M m0 = gtid < n ? sor_mask<REAL, VEC>(sp, e0) : (M)0;
		asm volatile("griddepcontrol.wait;\n" ::: "memory");
		for (uint32_t t = gtid; t < n; t += G) {
			const uint32_t e2 = t + 2u * G < n ? __ldg(list + t + 2u * G) : 0u;
			const M m1 = t + G < n ? sor_mask<REAL, VEC>(sp, e1) : (M)0;
			sor_vector_m<REAL, VEC, false>(sp, omega, one_minus_omega, e0, m0);
			e0 = e1; e1 = e2; m0 = m1;
		}
		return;
	}
	// Boundary blocks and interior blocks are disjoint: only the vectors of the slab's first and last plane read the
	// neighbours' ghost planes, so only the blocks that sweep them wait for the neighbours' epoch, and the system-scope
	// fence + signal that follow their stores stay off the interior blocks' path.
	// (with programmatic dependent launch: nothing below may run before the previous half sweep has completed)
	asm volatile("griddepcontrol.launch_dependents;\n" ::: "memory");
	asm volatile("griddepcontrol.wait;\n" ::: "memory");
	const uint32_t n_lo = __ldg(sh.n_lo), n_hi = __ldg(sh.n_hi), B = n_lo + (n - n_hi);
	const uint32_t nbb = min(max(gridDim.x / 4u, 1u), max(1u, (B + blockDim.x - 1) / blockDim.x));   // blocks that hold the boundary work
	if (blockIdx.x < nbb || gridDim.x == 1) {
		if (threadIdx.x == 0 && sh.peer_flag[0]) fs3_flag_wait(sh.my_flag, sh.wait_epoch, sh.flags, sh.nospin);
		if (threadIdx.x == 1 && sh.peer_flag[1]) fs3_flag_wait(sh.my_flag + 1, sh.wait_epoch, sh.flags, sh.nospin);
		__syncthreads();
		bool pushed = false;
		for (uint32_t t = gtid; t < B; t += nbb * blockDim.x) {
			sor_vector<REAL, VEC, true>(sp, omega, one_minus_omega, __ldg(list + (t < n_lo ? t : n_hi + (t - n_lo))));
			pushed = true;
		}
		if (pushed) __threadfence_system();   // my peer stores are visible system-wide before the block reports in
		__syncthreads();
		if (threadIdx.x == 0) {
			const uint32_t d = atomicAdd(sh.done, 1u);
			if (d == nbb - 1) {
				*sh.done = 0;
				__threadfence_system();   // ONE system-scope fence orders every block's (already fenced) stores before both flags
				if (sh.peer_flag[0]) fs3_st_relaxed_sys(sh.peer_flag[0], sh.signal_epoch);
				if (sh.peer_flag[1]) fs3_st_relaxed_sys(sh.peer_flag[1], sh.signal_epoch);
			}
		}
		if (gridDim.x > 1) return;
	}
	// interior: the remaining blocks (all of a one-block grid)
	const uint32_t ib = gridDim.x == 1 ? 0u : blockIdx.x - nbb, IG = gridDim.x == 1 ? blockDim.x : (gridDim.x - nbb) * blockDim.x;
	for (uint32_t t = ib * blockDim.x + threadIdx.x; t < n - B; t += IG) sor_vector<REAL, VEC, false>(sp, omega, one_minus_omega, __ldg(list + n_lo + t));
}

// ---- sharded solve in ONE cooperative launch, two-plane halos (multi-process handles, slabs >= 4 planes) ---------------------
// Launching every half sweep (round 1) cost 13 us each on 8 GPUs where the slab's memory traffic needs 3.5; one launch with
// one-plane halos pushed into the neighbours every half sweep (round 2, first version) still 13.8: time stamps taken inside that
// kernel (tools/sor_trace.py) showed that every half sweep carried  store to peer -> fence -> flag -> flight  and that
//   * a fence (gpu scope as well as system scope) of a CTA that has stores or atomics to a peer in flight waits for their
//     NVLink acknowledgements: 4 - 12 us;
//   * 148 remote atomics on ONE flag word serialise at the target: 9 us;
//   * sweeping entries whose operands a peer had just written took 4 us instead of 1;
//   * a CTA left the grid barrier (all CTAs polling the arrival word with acquire loads) 4 us after the last arrival.
// Hence this arrangement:
//   * TWO-plane halos, one exchange per ITERATION.  A rank also sweeps the RED cells of each neighbour's nearest plane itself
//     (the owner's b, diag, mask of that plane are copied once per projection, fs3i_sor_push_system; same inputs, same arithmetic:
//     bit-identical), so only BLACK values cross the boundary.  Lists: red = slab + one ghost plane per neighbour, black = slab.
//   * PULL.  The red EDGE entries (ghost plane + own first / last plane) read the neighbours' blacks straight from the
//     neighbours' arrays (ld.relaxed.sys, same offsets in every arena: sor_vector_m<EDGE>).  No sweeping CTA ever stores to a peer.
//   * Two flags per neighbour and iteration: B "my boundary blacks (first / last two planes) of iteration i are written", waited
//     for by the neighbour's red half sweep i + 1, and R "I have read your blacks of iteration i - 1", waited for by the
//     neighbour's black half sweep i before it overwrites them.  Dedicated boundary blocks sweep these entries FIRST, so both
//     flags are raised at the start of a half sweep and needed at the start of the next one.
//   * A COURIER CTA is the only one that writes to a peer: it waits on a local counter the boundary blocks bump, fences once at
//     system scope and stores the flag into both neighbours' arenas.  It takes no part in the grid barrier.
//   * Grid barrier of the sweeping CTAs: arrivals are atomics on one word; the last arriver publishes the generation in
//     FS3_SOR_GEN_LINES separate 128-byte lines (= L2 slices), block b polls copy b % LINES with relaxed loads and back-off: CTAs
//     polling ONE line while others still sweep take a large share of that slice's request rate.
//   * After the last iteration the boundary blocks fetch the neighbours' nearest black plane into the local ghost plane: the
//     gradient reads both colours of it.
// p travels through L2 only (ld.cg / write-through stores): L1 is not coherent across SMs inside a launch.
// sor_done (words): [0] boundary blocks that have reported, cumulative; [32] barrier arrivals; [32 r], r = 1..LINES, generation
// copies; [32 (LINES + 2 + parity)] ticket counters of the interior tail.  All zero at launch.
#define FS3_SOR_GEN_LINES 16
#define FS3_SOR_CHUNK 256   // entries per ticket of the dynamic tail of the interior sweep
#define FS3_SOR_MAX_NBB 256   // upper bound of the boundary blocks (a quarter of the grid is the effective one: 147 on a B200)
struct SorPersist {
	const void* peer_black[2];   // colour-1 array of the lower / upper neighbour (nullptr: no neighbour)
	uint32_t *peer_B[2], *peer_R[2];
	const uint32_t *my_B, *my_R;   // [0] written by the lower, [1] by the upper neighbour
	uint32_t e0, nbb;
	int iters;
	size_t lo_begin, hi_begin;     // first element of the slab's first / last plane in a colour array
	const uint32_t *n_lo[2], *n_hi[2], *list[2], *count[2];   // per colour: list entries [0, n_lo) and [n_hi, n) are the edge / boundary entries
	uint32_t *done, *gbar;
	int* flags;
#ifdef FS3_SOR_TRACE
	unsigned long long* trace;   // (tools/sor_trace.py)
#endif
};
#ifdef FS3_SOR_TRACE
__device__ __forceinline__ unsigned long long fs3_gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define SOR_TR(i) do { if (tr && threadIdx.x == 0) tr[i] = fs3_gtime(); } while (0)
#else
#define SOR_TR(i) do { } while (0)
#endif
template <typename REAL, int VEC>
__global__ void __launch_bounds__(FS3_SOR_THREADS, FS3_SOR_MINB) k_sor_persistent(Sys<REAL> sy, REAL omega, REAL one_minus_omega, SorPersist P) {
	const uint32_t G = gridDim.x - 1;   // sweeping blocks; the last block is the courier
	const uint32_t nbb = P.nbb;         // of which the first nbb sweep the edge / boundary entries
	if (blockIdx.x == G) {
		// ---- courier: the only CTA that stores to a neighbour.  When the boundary blocks of half sweep k have reported in (local
		// counter), one system-scope fence orders their stores before the flag it then writes into both neighbours' arenas.
		if (threadIdx.x == 0) {
			for (int k = 0; k < 2 * P.iters; k++) {
				const long long t0 = clock64();
				for (unsigned itn = 1;; itn++) {
					uint32_t v;
					asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(P.done) : "memory");
					if (v >= nbb * (uint32_t)(k + 1)) break;
					if ((itn & 1023u) == 0 && (*reinterpret_cast<volatile int*>(P.flags + 2) || clock64() - t0 > 8000000000LL)) { P.flags[2] = 1; return; }
				}
				asm volatile("fence.acq_rel.sys;\n" ::: "memory");
				uint32_t* const f0 = (k & 1) ? P.peer_B[0] : P.peer_R[0];
				uint32_t* const f1 = (k & 1) ? P.peer_B[1] : P.peer_R[1];
				if (f0) fs3_st_relaxed_sys(f0, P.e0 + (uint32_t)(k >> 1) + 1u);
				if (f1) fs3_st_relaxed_sys(f1, P.e0 + (uint32_t)(k >> 1) + 1u);
#ifdef FS3_SOR_TRACE
				if (k == 96 || k == 97) { unsigned long long* tr = P.trace + (64 + (size_t)(k - 96) * 1024 + blockIdx.x) * 8; tr[0] = tr[3] = fs3_gtime(); }
#endif
			}
		}
		return;
	}
	const bool bnd = blockIdx.x < nbb;
	__shared__ uint32_t s_chunk;
	SorShard sh;
	sh.peer_p[0] = const_cast<void*>(P.peer_black[0]); sh.peer_p[1] = const_cast<void*>(P.peer_black[1]);
	sh.lo_begin = P.lo_begin; sh.lo_end = 0; sh.hi_begin = P.hi_begin; sh.hi_end = 0;
	for (int k = 0; k < 2 * P.iters; k++) {
		const int c = k & 1, it = k >> 1;
#ifdef FS3_SOR_TRACE
		// half sweeps 64..95: first and last block; 96, 97: every block (rows of 8 after the first 32 * 2)
		unsigned long long* const tr = (k >= 64 && k < 96 && (blockIdx.x == 0 || blockIdx.x == G - 1)) ? P.trace + ((size_t)(k - 64) * 2 + (blockIdx.x == 0 ? 0 : 1)) * 8
		                               : (k == 96 || k == 97) ? P.trace + (64 + (size_t)(k - 96) * 1024 + blockIdx.x) * 8 : nullptr;
#endif
		SOR_TR(0);
		const SorPtrs<REAL> sp = sor_ptrs<REAL>(sy, c, sh);
		const uint32_t* __restrict__ list = c ? P.list[1] : P.list[0];
		const uint32_t n = __ldg(c ? P.count[1] : P.count[0]), n_lo = __ldg(c ? P.n_lo[1] : P.n_lo[0]), n_hi = __ldg(c ? P.n_hi[1] : P.n_hi[0]);
		const uint32_t B = n_lo + (n - n_hi);
		if (bnd) {
			// red waits for the neighbours' blacks of iteration it - 1 (flag B; none at it = 0: the host waited for "system set up"),
			// black for "the neighbour has read my blacks of iteration it - 1" (flag R of its red half sweep it)
			if (k > 0) {
				const uint32_t* const w = c ? P.my_R : P.my_B;
				if (threadIdx.x == 0 && P.peer_B[0]) fs3_flag_wait(w, P.e0 + (uint32_t)(it + c), P.flags, 0);
				if (threadIdx.x == 1 && P.peer_B[1]) fs3_flag_wait(w + 1, P.e0 + (uint32_t)(it + c), P.flags, 0);
			}
			__syncthreads();
			SOR_TR(1);
			{
				// (list two entries ahead, mask one ahead, as in the interior loop below: a thread's entries are serial)
				typedef typename VecT<REAL, VEC>::mtype M;
				const uint32_t BT = nbb * blockDim.x;
				uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
#define SOR_BENT(u) __ldg(list + ((u) < n_lo ? (u) : n_hi + ((u) - n_lo)))
				uint32_t e0 = t < B ? SOR_BENT(t) : 0u, e1 = t + BT < B ? SOR_BENT(t + BT) : 0u;
				M m0 = t < B ? sor_mask<REAL, VEC>(sp, e0) : (M)0;
				for (; t < B; t += BT) {
					const uint32_t e2 = t + 2u * BT < B ? SOR_BENT(t + 2u * BT) : 0u;
					const M m1 = t + BT < B ? sor_mask<REAL, VEC>(sp, e1) : (M)0;
					if (c) sor_vector_m<REAL, VEC, false, true>(sp, omega, one_minus_omega, e0, m0);
					else sor_vector_m<REAL, VEC, false, true, true>(sp, omega, one_minus_omega, e0, m0);
					e0 = e1; e1 = e2; m0 = m1;
				}
#undef SOR_BENT
			}
			SOR_TR(2);
			// report to the courier right away: the interior work below is not the neighbours' business
			__syncthreads();
			if (threadIdx.x == 0) { asm volatile("fence.acq_rel.gpu;\n" ::: "memory"); atomicAdd(P.done, 1u); }
		}
		// interior.  A static grid-stride split gives every thread floor or ceil of (entries / threads) entries -- 7 or 8 at 256^3
		// on two ranks: the barrier then waits 4 us for the blocks that got 8 -- and leaves the boundary blocks idle once they have
		// served the neighbours.  So all but the last two or three entries per interior thread are split statically (the same whole
		// number for every thread); the tail is handed out through a ticket counter, one entry per thread and ticket (a thread's
		// entries are serial, ~2 us each under load), to the boundary blocks as well.  Short sweeps (fewer than 4 entries per thread)
		// stay static: a ticket costs an L2 round trip.
		{
			const uint32_t nI = n - B, IT = (G - nbb) * blockDim.x, per = nI / IT;
			const uint32_t nS = per >= 4u ? (per - 2u) * IT : nI;
			if (!bnd) {
				// A thread's entries are serial and each is a chain of dependent loads (list -> mask -> p): the list entry is fetched
				// two entries ahead and the mask one ahead, so that only the p loads remain on the chain.
				typedef typename VecT<REAL, VEC>::mtype M;
				uint32_t t = (blockIdx.x - nbb) * blockDim.x + threadIdx.x;
				uint32_t e0 = t < nS ? __ldg(list + n_lo + t) : 0u, e1 = t + IT < nS ? __ldg(list + n_lo + t + IT) : 0u;
				M m0 = t < nS ? sor_mask<REAL, VEC>(sp, e0) : (M)0;
				for (; t < nS; t += IT) {
					const uint32_t e2 = t + 2u * IT < nS ? __ldg(list + n_lo + t + 2u * IT) : 0u;
					const M m1 = t + IT < nS ? sor_mask<REAL, VEC>(sp, e1) : (M)0;
					sor_vector_m<REAL, VEC, false, true>(sp, omega, one_minus_omega, e0, m0);
					e0 = e1; e1 = e2; m0 = m1;
				}
			}
			if (nS < nI) {
				uint32_t* const ticket = P.gbar + 32 * (FS3_SOR_GEN_LINES + 2 + (k & 1));
				const uint32_t nch = (nI - nS + FS3_SOR_CHUNK - 1) / FS3_SOR_CHUNK;
				if (threadIdx.x == 0) s_chunk = atomicAdd(ticket, 1u);
				__syncthreads();
				uint32_t ch = s_chunk;
				while (ch < nch) {
					__syncthreads();   // everybody has read s_chunk
					uint32_t nxt = 0;
					if (threadIdx.x == 0) nxt = atomicAdd(ticket, 1u);   // its latency overlaps the chunk's sweep
					for (uint32_t e = threadIdx.x; e < FS3_SOR_CHUNK; e += blockDim.x) {
						const uint32_t t = nS + ch * FS3_SOR_CHUNK + e;
						if (t < nI) sor_vector<REAL, VEC, false, true>(sp, omega, one_minus_omega, __ldg(list + n_lo + t));
					}
					if (threadIdx.x == 0) s_chunk = nxt;
					__syncthreads();
					ch = s_chunk;
				}
			}
		}
		// ---- boundary blocks report to the courier; grid barrier of the sweeping blocks ---------------------------------------------
		__syncthreads();
		SOR_TR(4);
		if (threadIdx.x == 0) {
			// (acq_rel fences: __threadfence() is the sequentially consistent one, which this protocol does not need)
			asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
			const uint32_t old = atomicAdd(P.gbar, 1u);
			if (old == (uint32_t)(k + 1) * G - 1u) {
				asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
				// the ticket counter of the next half sweep's parity was last used two half sweeps ago: reset it before anybody leaves
				asm volatile("st.relaxed.gpu.global.u32 [%0], %1;\n" ::"l"(P.gbar + 32 * (FS3_SOR_GEN_LINES + 2 + ((k + 1) & 1))), "r"(0u) : "memory");
				asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
				for (int r = 1; r <= FS3_SOR_GEN_LINES; r++) asm volatile("st.relaxed.gpu.global.u32 [%0], %1;\n" ::"l"(P.gbar + 32 * r), "r"((uint32_t)k + 1u) : "memory");
				SOR_TR(5);
			} else {
				SOR_TR(5);
				const uint32_t* const gen = P.gbar + 32 * (1 + (blockIdx.x % FS3_SOR_GEN_LINES));
				const long long t0 = clock64();
				for (unsigned itn = 1;; itn++) {
					uint32_t v;
					asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(gen) : "memory");
					if (v >= (uint32_t)k + 1u) break;
					__nanosleep(100);
					if ((itn & 1023u) == 0 && (*reinterpret_cast<volatile int*>(P.flags + 2) || clock64() - t0 > 8000000000LL)) { P.flags[2] = 1; break; }
				}
				asm volatile("fence.acq_rel.gpu;\n" ::: "memory");
			}
		}
		__syncthreads();
		SOR_TR(6);
	}
	// ---- the gradient reads the neighbours' nearest plane of both colours from the local ghost planes: red is there (swept here),
	// black is fetched now, once the neighbours' last black half sweep has been flagged
	if (bnd) {
		if (threadIdx.x == 0 && P.peer_B[0]) fs3_flag_wait(P.my_B, P.e0 + (uint32_t)P.iters, P.flags, 0);
		if (threadIdx.x == 1 && P.peer_B[1]) fs3_flag_wait(P.my_B + 1, P.e0 + (uint32_t)P.iters, P.flags, 0);
		__syncthreads();
		REAL* const black = sy.p[1];
		const size_t pl = sy.rb.plane;
		for (int side = 0; side < 2; side++) {
			const REAL* const src = reinterpret_cast<const REAL*>(side ? P.peer_black[1] : P.peer_black[0]);
			if (!src) continue;
			const size_t a = side ? P.hi_begin + pl : P.lo_begin - pl;
			for (size_t v = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * VEC; v < pl; v += (size_t)nbb * blockDim.x * VEC) {
				REAL t[VEC];
				ldv_sys(src + a + v, t);
				stv(black + a + v, t);
			}
		}
	}
}

// active-vector flags (+1 trailing zero so that the exclusive scan's last entry is the count) and compaction
template <int VEC>
__global__ void __launch_bounds__(256) k_sor_flags(const unsigned char* __restrict__ mask, size_t nvec, uint32_t* __restrict__ flags, size_t v0, size_t v1) {
	size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (v > nvec) return;
	uint32_t f = 0;
	if (v >= v0 && v < v1) {
		if (VEC == 2) f = (*reinterpret_cast<const unsigned short*>(mask + v * 2) & 0x4040u) != 0;
		else f = (*reinterpret_cast<const unsigned int*>(mask + v * 4) & 0x40404040u) != 0;
	}
	flags[v] = f;
}
template <int VEC>
__global__ void __launch_bounds__(256) k_sor_compact(const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos, size_t nvec, RBLayout rb,
                                                     int colour, uint32_t* __restrict__ list) {
	size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (v >= nvec || !flags[v]) return;
	size_t row = (v * VEC) / rb.hp;   // = y + ny*(z+1)
	int y = (int)(row % rb.ny), z = (int)(row / rb.ny) - 1;
	list[pos[v]] = (uint32_t)v | ((uint32_t)((colour + y + z) & 1) << 31);
}

// ---- residual ||b - A p|| over fluid cells (instrumentation; not in the reference) ------------------
template <typename REAL>
__global__ void __launch_bounds__(256) k_sor_residual(Sys<REAL> sy, double* __restrict__ out) {
	const int nx = sy.rb.nx, ny = sy.rb.ny, nz = sy.rb.nz;
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	double r2 = 0, ra = 0;
	if (c < (size_t)nx * ny * nz) {
		int x = (int)(c % nx), y = (int)((c / nx) % ny), z = (int)(c / ((size_t)nx * ny));
		int col = (x + y + z) & 1, s = x & 1, xh = x >> 1;
		size_t q = sy.rb.at(xh, y, z);
		unsigned m = sy.mask[col][q];
		if (m & 64u) {
			const REAL* o = sy.p[col ^ 1];
			double nms = 0;
			if (m & 1u) nms -= (double)o[q - 1 + s];
			if (m & 2u) nms -= (double)o[q + s];
			if (m & 4u) nms -= (double)o[q - sy.rb.hp];
			if (m & 8u) nms -= (double)o[q + sy.rb.hp];
			if (m & 16u) nms -= (double)o[q - sy.rb.plane];
			if (m & 32u) nms -= (double)o[q + sy.rb.plane];
			double r = (double)sy.b[col][q] - ((double)sy.dg[col][q] * (double)sy.p[col][q] + nms);
			r2 = r * r;
			ra = fabs(r);
		}
	}
	__shared__ double s2[8], sa[8];
	for (int o = 16; o; o >>= 1) { r2 += __shfl_down_sync(0xffffffffu, r2, o); ra = fmax(ra, __shfl_down_sync(0xffffffffu, ra, o)); }
	int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	if (lane == 0) { s2[w] = r2; sa[w] = ra; }
	__syncthreads();
	if (threadIdx.x == 0) {
		for (int k = 1; k < 8; k++) { r2 += s2[k]; ra = fmax(ra, sa[k]); }
		atomicAdd(out, r2);
		atomicMax(reinterpret_cast<unsigned long long*>(out + 1), (unsigned long long)__double_as_longlong(ra));
	}
}

// ---- III. pressure-gradient update, four ghost-fluid cases per face (S3D.cpp:1012-1084) -------------
template <typename REAL>
__device__ __forceinline__ float grad_face(float g, float plo, float phi_hi, REAL p_lo, REAL p_hi, REAL scale, REAL max_ratio) {
	typedef Ops<REAL> O;
	REAL lo = (REAL)plo, hi = (REAL)phi_hi;
	if (lo < 0 && hi < 0) return (float)O::sub((REAL)g, O::mul(scale, O::sub(p_hi, p_lo)));
	if (lo < 0 && hi >= 0) {
		REAL r = O::div(-hi, lo); r = r < (REAL)0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);
		return (float)O::add((REAL)g, O::mul(O::mul(scale, O::add((REAL)1, r)), p_lo));
	}
	if (lo >= 0 && hi < 0) {
		REAL r = O::div(-lo, hi); r = r < (REAL)0 ? (REAL)0 : (r > max_ratio ? max_ratio : r);
		return (float)O::sub((REAL)g, O::mul(O::mul(scale, O::add((REAL)1, r)), p_hi));
	}
	return 0.0f;
}

template <typename REAL>
__global__ void __launch_bounds__(256) k_project_gradient(Sys<REAL> sy, const float* __restrict__ sU, const float* __restrict__ sV,
                                                          const float* __restrict__ sW, const float* __restrict__ phi, float* __restrict__ dU,
                                                          float* __restrict__ dV, float* __restrict__ dW, REAL scale, float gdt, int add_g,
                                                          REAL max_ratio, size_t c0, size_t c1) {
	const int nx = sy.rb.nx, ny = sy.rb.ny, nz = sy.rb.nz;
	size_t c = c0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= c1) return;
	int x = (int)(c % nx), y = (int)((c / nx) % ny), z = (int)(c / ((size_t)nx * ny));
	int col = (x + y + z) & 1, s = x & 1;
	size_t q = sy.rb.at(x >> 1, y, z);
	REAL pc = sy.p[col][q];
	float fc = phi[c];
	const REAL* o = sy.p[col ^ 1];
	size_t iu = (size_t)x + (size_t)(nx + 1) * ((size_t)y + (size_t)ny * z);
	size_t iv = (size_t)x + (size_t)nx * ((size_t)y + (size_t)(ny + 1) * z);
	size_t iw = c, sZ = (size_t)nx * ny;
	// x face above this cell; wall faces are zero (SetEdgeVelocitiesToZero, S3D.cpp:1012)
	if (x == 0) dU[iu] = 0.f;
	dU[iu + 1] = x == nx - 1 ? 0.f : grad_face<REAL>(sU[iu + 1], fc, phi[c + 1], pc, o[q + s], scale, max_ratio);
	if (y == 0) dV[iv] = 0.f;
	{
		float g = sV[iv + nx];
		if (add_g) g = __fadd_rn(g, gdt);
		dV[iv + nx] = y == ny - 1 ? 0.f : grad_face<REAL>(g, fc, phi[c + nx], pc, o[q + sy.rb.hp], scale, max_ratio);
	}
	if (z == 0) dW[iw] = 0.f;
	dW[iw + sZ] = z == nz - 1 ? 0.f : grad_face<REAL>(sW[iw + sZ], fc, phi[c + sZ], pc, o[q + sy.rb.plane], scale, max_ratio);
}

template <typename REAL>
__global__ void k_export_system(Sys<REAL> sy, double* __restrict__ p, double* __restrict__ b, double* __restrict__ dg) {
	const int nx = sy.rb.nx, ny = sy.rb.ny, nz = sy.rb.nz;
	size_t c = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (c >= (size_t)nx * ny * nz) return;
	int x = (int)(c % nx), y = (int)((c / nx) % ny), z = (int)(c / ((size_t)nx * ny));
	int col = (x + y + z) & 1;
	size_t q = sy.rb.at(x >> 1, y, z);
	if (p) p[c] = (double)sy.p[col][q];
	if (b) b[c] = (double)sy.b[col][q];
	if (dg) dg[c] = (double)sy.dg[col][q];
}

template <typename REAL>
__global__ void k_clear_p(Sys<REAL> sy) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < sy.rb.total) { sy.p[0][q] = 0; sy.p[1][q] = 0; }
}

// ---- host side --------------------------------------------------------------------------------------
template <typename REAL> static Sys<REAL> make_sys(fs3_sim* s) {
	Sys<REAL> sy;
	for (int c = 0; c < 2; c++) {
		sy.p[c] = (REAL*)s->p[c]; sy.b[c] = (REAL*)s->b[c]; sy.dg[c] = (REAL*)s->dg[c]; sy.mask[c] = s->mask[c];
	}
	sy.rb = s->rb;
	return sy;
}

int fs3i_alloc_pressure(fs3_sim* s) {
	int prec = s->prm.precision ? 1 : 0;
	if (s->p[0] && s->alloc_precision == prec) return FS3_OK;
	if (s->p[0] && s->world > 1) {
		// a sharded handle allocates the fp64-sized system once (arena offsets must agree on all ranks) and reinterprets it
		s->alloc_precision = prec; s->system_valid = false; s->sor_list_valid = false;
		return FS3_OK;
	}
	for (int c = 0; c < 2; c++) {
		fs3i_free(s, s->p[c]); fs3i_free(s, s->b[c]); fs3i_free(s, s->dg[c]); fs3i_free(s, s->mask[c]);
		fs3i_free(s, s->sor_list[c]); fs3i_free(s, s->sor_pos[c]);
		s->p[c] = s->b[c] = s->dg[c] = nullptr; s->mask[c] = nullptr; s->sor_list[c] = s->sor_pos[c] = nullptr;
	}
	fs3i_free(s, s->sor_flags);
	s->sor_flags = nullptr;
	RBLayout& rb = s->rb;
	rb.nx = s->nx; rb.ny = s->ny; rb.nz = s->nz;
	rb.hx = (s->nx + 1) / 2;
	rb.hp = (rb.hx + 3) / 4 * 4;
	rb.plane = (size_t)rb.hp * s->ny;
	rb.total = rb.plane * (size_t)(s->nz + 2);
	size_t es = (prec && s->world == 1) ? 4 : 8;
	int rc;
	for (int c = 0; c < 2; c++) {
		if ((rc = fs3i_alloc(s, &s->p[c], rb.total * es)) || (rc = fs3i_alloc(s, &s->b[c], rb.total * es)) ||
		    (rc = fs3i_alloc(s, &s->dg[c], rb.total * es)) || (rc = fs3i_alloc(s, (void**)&s->mask[c], rb.total))) return rc;
		FS3_CUDA(cudaMemsetAsync(s->p[c], 0, rb.total * es, s->stream)); FS3_CUDA(cudaMemsetAsync(s->b[c], 0, rb.total * es, s->stream));
		FS3_CUDA(cudaMemsetAsync(s->dg[c], 0, rb.total * es, s->stream)); FS3_CUDA(cudaMemsetAsync(s->mask[c], 0, rb.total, s->stream));
	}
	const size_t nvec = rb.total / ((prec && s->world == 1) ? 4 : 2);
	if ((rc = fs3i_alloc(s, (void**)&s->sor_flags, (nvec + 1) * sizeof(uint32_t)))) return rc;
	for (int c = 0; c < 2; c++) {
		if ((rc = fs3i_alloc(s, (void**)&s->sor_list[c], nvec * sizeof(uint32_t)))) return rc;
		if ((rc = fs3i_alloc(s, (void**)&s->sor_pos[c], (nvec + 1) * sizeof(uint32_t)))) return rc;
	}
	if ((nvec + 1) / 4096 + 2 > s->scan_block_cap) { s->err = "scan scratch too small for the SOR lists"; return FS3_ERR_STATE; }
	s->alloc_precision = prec;
	s->system_valid = false;
	s->sor_list_valid = false;
	return FS3_OK;
}

static double fs3_omega(const fs3_sim* s) { return s->prm.omega > 0 ? s->prm.omega : 2 - 3.16343 / s->nx; }  // S3D.cpp:944

// active-vector lists of both colours (flags -> exclusive scan -> compaction) over this rank's slab; counts stay on the device
template <typename REAL, int VEC> static int sor_prepare_t(fs3_sim* s) {
	if (s->sor_list_valid) return FS3_OK;
	const size_t nvec = s->rb.total / VEC;
	for (int c = 0; c < 2; c++) {
		// two-plane halos (fs3i_sor_two_plane): the red list also holds the neighbours' nearest plane
		const int g = (c == 0 && fs3i_sor_two_plane(s)) ? 1 : 0;
		const int za = s->z0 - ((g && s->rank > 0) ? 1 : 0), zb = s->z1 + ((g && s->rank < s->world - 1) ? 1 : 0);
		const size_t v0 = s->rb.plane * (size_t)(za + 1) / VEC, v1 = s->rb.plane * (size_t)(zb + 1) / VEC;
		k_sor_flags<VEC><<<cdiv((long long)nvec + 1, 256), 256, 0, s->stream>>>(s->mask[c], nvec, s->sor_flags, v0, v1); FS3_LAUNCH_CHECK();
		int rc = fs3i_scan_u32(s, s->sor_flags, s->sor_pos[c], nvec + 1);
		if (rc) return rc;
		k_sor_compact<VEC><<<cdiv((long long)nvec, 256), 256, 0, s->stream>>>(s->sor_flags, s->sor_pos[c], nvec, s->rb, c, s->sor_list[c]); FS3_LAUNCH_CHECK();
		s->sor_count[c] = s->sor_pos[c] + nvec;
	}
	s->sor_list_valid = true;
	return FS3_OK;
}
int fs3i_sor_prepare(fs3_sim* s) { return s->alloc_precision ? sor_prepare_t<float, 4>(s) : sor_prepare_t<double, 2>(s); }

// L2 residency of p during the solve.  The two colours of p are the only operands one half sweep shares with the next; b,
// the mask and the list stream through once per sweep and, left alone, keep pushing p out of the 126 MB L2.  For the length
// of fs3i_sor_loop a persisting L2 carve-out is raised (it is a real partition: while it exists every other kernel has
// that much less L2, hence the stream synchronisation + cudaDeviceSetLimit on both sides of the solve) and every half
// sweep is launched with an access-policy window over the colour it writes.  The fraction of the window that persists
// is carve-out / bytes of p the active vectors touch, so that the persisting lines never thrash among themselves; below
// 0.5 (large grids) or for grids that fit in L2 anyway the feature stays off and the solve runs without the two syncs.
static size_t sor_persist_cap(fs3_sim* s) {
	if (s->persist_cap < 0) {   // per handle (= per device): the largest carve-out the device offers
		cudaDeviceProp pr;
		s->persist_cap = 0;
		if (cudaGetDeviceProperties(&pr, s->device) == cudaSuccess && (size_t)pr.accessPolicyMaxWindowSize >= ((size_t)64 << 20))
			s->persist_cap = (long long)pr.persistingL2CacheMaxSize;
	}
	return (size_t)s->persist_cap;
}
// half sweep k (colour k & 1) of the current projection
template <typename REAL, int VEC> static int sor_half_t(fs3_sim* s, int k) {
	Sys<REAL> sy = make_sys<REAL>(s);
	REAL omega = (REAL)fs3_omega(s);
	if (sizeof(REAL) == 4) omega = (REAL)2 - (REAL)3.16343 / s->nx, omega = s->prm.omega > 0 ? (REAL)s->prm.omega : omega;
	REAL omo = (REAL)1 - omega;
	// persistent grid-stride launch: enough blocks to fill the machine; the list length is read on the device
	size_t threads = (size_t)(s->rb.hp / VEC) * s->ny * (s->z1 - s->z0);
	if (!s->sor_per_sm) {    // resident blocks per SM of the sweep kernel: the grid is exactly one wave
		int a = 0, b = 0;
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, k_sor_half<REAL, VEC, false>, FS3_SOR_THREADS, 0);
		cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, k_sor_half<REAL, VEC, true>, FS3_SOR_THREADS, 0);
		s->sor_per_sm = std::max(1, std::min(a, b));
	}
	const int per_sm = s->sor_per_sm;
	int nb = (int)std::min<long long>(std::max(1, cdiv((long long)threads, FS3_SOR_THREADS)), (long long)s->sm_count * per_sm);
	const int c = k & 1;
	if (s->world == 1) {
		// half sweeps after the first are launched with programmatic stream serialization (see the kernel); the first one
		// follows the set-up kernels with a normal dependency
		cudaLaunchConfig_t cfg = {};
		cfg.gridDim = dim3(nb); cfg.blockDim = dim3(FS3_SOR_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = s->stream;
		cudaLaunchAttribute at[2];
		int na = 0;
		if (s->sor_persist_ratio > 0) {
			at[na].id = cudaLaunchAttributeAccessPolicyWindow;
			at[na].val.accessPolicyWindow.base_ptr = s->p[c];
			at[na].val.accessPolicyWindow.num_bytes = s->rb.total * sizeof(REAL);
			at[na].val.accessPolicyWindow.hitRatio = s->sor_persist_ratio;
			at[na].val.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
			at[na].val.accessPolicyWindow.missProp = cudaAccessPropertyNormal;
			na++;
		}
		if (k > 0 && !s->prm.record_residuals) {
			at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
			at[na].val.programmaticStreamSerializationAllowed = 1;
			na++;
		}
		cfg.attrs = at; cfg.numAttrs = na;
		const uint32_t* lst = s->sor_list[c]; const uint32_t* cnt = s->sor_count[c];
		FS3_CUDA(cudaLaunchKernelEx(&cfg, k_sor_half<REAL, VEC, false>, sy, c, omega, omo, lst, cnt, SorShard{}));
		FS3_LAUNCH_CHECK();
		return FS3_OK;
	}
	SorShard sh;
	const bool lo = s->rank > 0, hi = s->rank < s->world - 1;
	sh.peer_p[0] = lo ? fs3_peer(s, s->rank - 1, (char*)s->p[c]) : nullptr;
	sh.peer_p[1] = hi ? fs3_peer(s, s->rank + 1, (char*)s->p[c]) : nullptr;
	// my epoch goes to the lower neighbour's "from above" slot and to the upper neighbour's "from below" slot
	sh.peer_flag[0] = lo ? fs3_peer(s, s->rank - 1, s->bar) + FS3_MAX_WORLD + 1 : nullptr;
	sh.peer_flag[1] = hi ? fs3_peer(s, s->rank + 1, s->bar) + FS3_MAX_WORLD + 0 : nullptr;
	sh.my_flag = s->bar + FS3_MAX_WORLD;
	sh.wait_epoch = s->sor_e0 + (uint32_t)k;
	sh.signal_epoch = s->sor_e0 + (uint32_t)k + 1u;
	sh.lo_begin = s->rb.plane * (size_t)(s->z0 + 1); sh.lo_end = sh.lo_begin + s->rb.plane;
	sh.hi_begin = s->rb.plane * (size_t)s->z1; sh.hi_end = sh.hi_begin + s->rb.plane;
	sh.done = s->sor_done; sh.flags = s->dev_flags; sh.nospin = s->lockstep ? 1 : 0;
	// exclusive scan of the activity flags at the end of the first plane / the start of the last plane of the slab
	sh.n_lo = s->sor_pos[c] + s->rb.plane * (size_t)(s->z0 + 2) / VEC;
	sh.n_hi = s->sor_pos[c] + s->rb.plane * (size_t)s->z1 / VEC;
	{
		// (programmatic dependent launch stays off here: measured 12.7 -> 14.9 us per half sweep on 8 GPUs, the boundary
		// blocks' flag wait then holds SM slots the previous sweep's interior blocks need)
		cudaLaunchConfig_t cfg = {};
		cfg.gridDim = dim3(nb); cfg.blockDim = dim3(FS3_SOR_THREADS); cfg.dynamicSmemBytes = 0; cfg.stream = s->stream;
		cfg.attrs = nullptr; cfg.numAttrs = 0;
		const uint32_t* lst = s->sor_list[c]; const uint32_t* cnt = s->sor_count[c];
		FS3_CUDA(cudaLaunchKernelEx(&cfg, k_sor_half<REAL, VEC, true>, sy, c, omega, omo, lst, cnt, sh));
		FS3_LAUNCH_CHECK();
	}
	s->sor_epoch = sh.signal_epoch;
	return FS3_OK;
}
int fs3i_sor_half_launch(fs3_sim* s, int k) { return s->alloc_precision ? sor_half_t<float, 4>(s, k) : sor_half_t<double, 2>(s, k); }

// two-plane halos: multi-process handles whose slab has room for two distinct boundary planes on either side
bool fs3i_sor_two_plane(const fs3_sim* s) { return s->world > 1 && !s->lockstep && s->nz / s->world >= 4; }

template <typename REAL, int VEC> static int sor_persistent_t(fs3_sim* s, int iters) {
	Sys<REAL> sy = make_sys<REAL>(s);
	REAL omega = (REAL)fs3_omega(s);
	if (sizeof(REAL) == 4) omega = (REAL)2 - (REAL)3.16343 / s->nx, omega = s->prm.omega > 0 ? (REAL)s->prm.omega : omega;
	REAL omo = (REAL)1 - omega;
	int per_sm = 0;
	FS3_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_sor_persistent<REAL, VEC>, FS3_SOR_THREADS, 0));
	per_sm = std::max(1, per_sm);
	const size_t threads = (size_t)(s->rb.hp / VEC) * s->ny * (s->z1 - s->z0 + 2);
	// sweeping blocks + the courier, all co-resident
	const int nb = (int)std::min<long long>(std::max<long long>(2, cdiv((long long)threads, FS3_SOR_THREADS)), (long long)s->sm_count * per_sm - 1);
	if (nb < 2) { s->err = "device too small for the one-launch sharded solve"; return FS3_ERR_STATE; }
	SorPersist P;
	const bool lo = s->rank > 0, hi = s->rank < s->world - 1;
	const size_t pl = s->rb.plane;
	uint32_t* const bar_lo = lo ? fs3_peer(s, s->rank - 1, s->bar) + FS3_MAX_WORLD : nullptr;
	uint32_t* const bar_hi = hi ? fs3_peer(s, s->rank + 1, s->bar) + FS3_MAX_WORLD : nullptr;
	P.peer_black[0] = lo ? fs3_peer(s, s->rank - 1, (char*)s->p[1]) : nullptr;
	P.peer_black[1] = hi ? fs3_peer(s, s->rank + 1, (char*)s->p[1]) : nullptr;
	// my flags go to the lower neighbour's "from above" slots (+1, +3) and to the upper neighbour's "from below" slots (+0, +2)
	P.peer_B[0] = lo ? bar_lo + 1 : nullptr; P.peer_B[1] = hi ? bar_hi + 0 : nullptr;
	P.peer_R[0] = lo ? bar_lo + 3 : nullptr; P.peer_R[1] = hi ? bar_hi + 2 : nullptr;
	P.my_B = s->bar + FS3_MAX_WORLD; P.my_R = s->bar + FS3_MAX_WORLD + 2;
	P.e0 = s->sor_e0; P.iters = iters;
	// Boundary blocks do little else (they wait on the neighbours, and their entries read remote memory), and the sweep's throughput
	// is proportional to the blocks that work on it, so their number follows their share of the planes (two per neighbour; x 3 for
	// the waits and the slower entries): 256^3: 147 of 591 blocks on an inner rank of 8, 107 of 4, 27 at 2 ranks.
	const int edge_planes = 2 * ((lo ? 1 : 0) + (hi ? 1 : 0));
	P.nbb = (uint32_t)std::min(std::min(std::max(nb / 4, 1), FS3_SOR_MAX_NBB), std::max(8, (int)(3LL * nb * edge_planes / (s->z1 - s->z0 + 2))));
	P.nbb = std::min<uint32_t>(P.nbb, (uint32_t)nb - 1u);
	// array plane index of cell plane z is z + 1
	P.lo_begin = pl * (size_t)(s->z0 + 1); P.hi_begin = pl * (size_t)s->z1;
	// red: planes z0-1, z0 and z1-1, z1 read the neighbours' blacks; black: planes z0, z0+1 and z1-2, z1-1 are what the neighbours read.
	// Without a neighbour the range is empty (the scan is 0 up to the list's first plane and n from its last one on).
	P.n_lo[0] = s->sor_pos[0] + pl * (size_t)(lo ? s->z0 + 2 : s->z0 + 1) / VEC;
	P.n_hi[0] = s->sor_pos[0] + pl * (size_t)(hi ? s->z1 : s->z1 + 1) / VEC;
	P.n_lo[1] = s->sor_pos[1] + pl * (size_t)(lo ? s->z0 + 3 : s->z0 + 1) / VEC;
	P.n_hi[1] = s->sor_pos[1] + pl * (size_t)(hi ? s->z1 - 1 : s->z1 + 1) / VEC;
	for (int c = 0; c < 2; c++) { P.list[c] = s->sor_list[c]; P.count[c] = s->sor_count[c]; }
	P.done = s->sor_done; P.gbar = s->sor_done + 32; P.flags = s->dev_flags;
	FS3_CUDA(cudaMemsetAsync(s->sor_done, 0, 4096, s->stream));
#ifdef FS3_SOR_TRACE
	static unsigned long long* trace_buf = nullptr;
	static int trace_calls = 0;
	if (!trace_buf) FS3_CUDA(cudaMallocManaged((void**)&trace_buf, (64 + 2048) * 8 * sizeof(unsigned long long)));
	P.trace = trace_buf;
#endif
	void* args[] = {(void*)&sy, (void*)&omega, (void*)&omo, (void*)&P};
	FS3_CUDA(cudaLaunchCooperativeKernel((const void*)k_sor_persistent<REAL, VEC>, dim3(nb + 1), dim3(FS3_SOR_THREADS), args, 0, s->stream));
	FS3_LAUNCH_CHECK();
#ifdef FS3_SOR_TRACE
	if (++trace_calls == 4) {
		FS3_CUDA(cudaStreamSynchronize(s->stream));
		std::string out;
		char buf[256];
		for (int k = 0; k < 32; k++)
			for (int b = 0; b < 2; b++) {
				const unsigned long long* t = trace_buf + ((size_t)k * 2 + b) * 8;
				snprintf(buf, sizeof buf, "sortrace rank %d half %d %s", s->rank, k + 64, b ? "interior" : "boundary"); out += buf;
				for (int i = 1; i < 7; i++) { snprintf(buf, sizeof buf, " %6lld", t[i] >= t[0] && t[i] ? (long long)(t[i] - t[0]) : -1LL); out += buf; }
				snprintf(buf, sizeof buf, "  start +%lld\n", k ? (long long)(t[0] - trace_buf[((size_t)(k - 1) * 2 + b) * 8]) : 0LL); out += buf;
			}
		for (int h = 0; h < 2; h++) {
			const unsigned long long* base = trace_buf + (64 + (size_t)h * 1024) * 8;
			unsigned long long tmin = ~0ULL;
			for (int b = 0; b < nb; b++) tmin = std::min(tmin, base[(size_t)b * 8]);
			for (int b = 0; b <= nb; b++) {
				const unsigned long long* t = base + (size_t)b * 8;
				snprintf(buf, sizeof buf, "sorblocks rank %d half %d block %d%s", s->rank, 96 + h, b, b < (int)P.nbb ? " B" : ""); out += buf;
				for (int i = 0; i < 7; i++) { snprintf(buf, sizeof buf, " %6lld", t[i] ? (long long)(t[i] - tmin) : -1LL); out += buf; }
				out += "\n";
			}
		}
		snprintf(buf, sizeof buf, "gpurun_out/sortrace_rank%d.txt", s->rank);
		if (FILE* f = fopen(buf, "w")) { fwrite(out.data(), 1, out.size(), f); fclose(f); }
	}
#endif
	s->sor_epoch = s->sor_e0 + 2u * (uint32_t)iters;   // (the flags end at e0 + iters; the next projection's e0 lies above either scheme's range)
	return FS3_OK;
}
int fs3i_sor_persistent(fs3_sim* s, int iters) {
	if (!fs3i_sor_two_plane(s)) { s->err = "the one-launch solve needs a multi-process handle with slabs of at least 4 planes"; return FS3_ERR_STATE; }
	return s->alloc_precision ? sor_persistent_t<float, 4>(s, iters) : sor_persistent_t<double, 2>(s, iters);
}
// two-plane halos: the system (b, diag, mask) of the RED cells of my first / last plane goes to the neighbour that sweeps them too
int fs3i_sor_push_system(fs3_sim* s) {
	if (!fs3i_sor_two_plane(s)) return FS3_OK;
	const size_t es = s->alloc_precision ? 4 : 8, pl = s->rb.plane;
	for (int side = 0; side < 2; side++) {
		const int r = side ? s->rank + 1 : s->rank - 1;
		if (r < 0 || r >= s->world) continue;
		const size_t off = pl * (size_t)(side ? s->z1 : s->z0 + 1);   // array plane of cell plane z1 - 1 / z0
		FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, r, (char*)s->b[0]) + off * es, (char*)s->b[0] + off * es, pl * es, cudaMemcpyDefault, s->stream));
		FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, r, (char*)s->dg[0]) + off * es, (char*)s->dg[0] + off * es, pl * es, cudaMemcpyDefault, s->stream));
		FS3_CUDA(cudaMemcpyAsync(fs3_peer(s, r, (char*)s->mask[0]) + off, (char*)s->mask[0] + off, pl, cudaMemcpyDefault, s->stream));
	}
	return FS3_OK;
}

// The persisting-L2 size is a DEVICE-wide limit: whatever the host application (or another handle) had configured is put
// back when the solve ends, on every path out of fs3i_sor_loop.
struct PersistLimitGuard {
	size_t prev = 0;
	bool raised = false;
	bool raise(size_t bytes) {
		if (cudaDeviceGetLimit(&prev, cudaLimitPersistingL2CacheSize) != cudaSuccess) return false;
		if (prev >= bytes) return true;   // the application already keeps a carve-out at least this large: use it, leave it
		if (cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, bytes) != cudaSuccess) { cudaGetLastError(); return false; }
		raised = true;
		return true;
	}
	void restore(cudaStream_t st) {
		if (!raised) return;
		cudaStreamSynchronize(st);
		cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, prev);
		raised = false;
	}
};

int fs3i_sor_loop(fs3_sim* s, int iters) {
	int rc = fs3i_sor_prepare(s);
	if (rc) return rc;
	PersistLimitGuard guard;
	bool rec = s->prm.record_residuals != 0;
	if (rec) {
		if (s->world > 1) { s->err = "residual recording is not available on a sharded handle"; return FS3_ERR_STATE; }
		if (iters > s->res_cap) { s->err = "residual buffer too small"; return FS3_ERR_STATE; }
		FS3_CUDA(cudaMemsetAsync(s->res, 0, sizeof(double) * 2 * iters, s->stream));
	}
	s->sor_halves = 2 * iters;
	s->sor_persist_ratio = 0;
	const size_t es = s->alloc_precision ? 4 : 8, cap = s->world == 1 && !rec ? sor_persist_cap(s) : 0;
	if (cap && 2 * s->rb.total * es > ((size_t)48 << 20) && s->rb.total * es <= ((size_t)128 << 20)) {
		uint32_t cnt[2];
		FS3_CUDA(cudaMemcpyAsync(&cnt[0], s->sor_count[0], 4, cudaMemcpyDeviceToHost, s->stream));
		FS3_CUDA(cudaMemcpyAsync(&cnt[1], s->sor_count[1], 4, cudaMemcpyDeviceToHost, s->stream));
		FS3_CUDA(cudaStreamSynchronize(s->stream));
		const double touched = ((double)cnt[0] + cnt[1]) * 16.0;   // one 128-bit vector per list entry
		const double ratio = touched > 0 ? std::min(1.0, 0.95 * (double)cap / touched) : 0.0;
		if (ratio >= 0.5 && touched > (double)((size_t)32 << 20) && guard.raise(cap)) s->sor_persist_ratio = (float)ratio;
	}
	for (int it = 0; it < iters && !rc; it++) {
		if ((rc = fs3i_sor_half_launch(s, 2 * it)) || (rc = fs3i_sor_half_launch(s, 2 * it + 1))) break;
		if (rec) {
			if (s->alloc_precision) k_sor_residual<float><<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(make_sys<float>(s), s->res + 2 * it);
			else k_sor_residual<double><<<cdiv((long long)s->nC, 256), 256, 0, s->stream>>>(make_sys<double>(s), s->res + 2 * it);
			s->launches++;
			if (cudaGetLastError() != cudaSuccess) { s->err = "residual kernel launch failed"; rc = FS3_ERR_CUDA; }
		}
	}
	guard.restore(s->stream);   // also on the error paths above
	s->sor_persist_ratio = 0;
	if (rc) return rc;
	s->res_count = rec ? iters : 0;
	return FS3_OK;
}

template <typename REAL> static int project_setup_t(fs3_sim* s, float dt, bool from_old, bool with_gravity) {
	Sys<REAL> sy = make_sys<REAL>(s);
	const float *sU = from_old ? s->oU : s->U, *sV = from_old ? s->oV : s->V, *sW = from_old ? s->oW : s->W;
	REAL dx = (REAL)1.0 / s->cpm;                       // S3D.cpp:834
	REAL scale = -dx * s->prm.rho / dt;                 // S3D.cpp:835
	float gdt = s->prm.gravity_y * dt;                  // S3D.cpp:784
	const size_t plane = (size_t)s->nx * s->ny, c0 = plane * s->z0, c1 = plane * s->z1;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_PROJECT_SETUP], s->stream);
	k_project_setup<REAL><<<cdiv((long long)(c1 - c0), 256), 256, 0, s->stream>>>(sy, sU, sV, sW, s->phi, scale, gdt, with_gravity ? 1 : 0, (REAL)1000.0, c0, c1);
	FS3_LAUNCH_CHECK();
	if (s->world > 1) {
		// p = 0 also on the ghost planes the neighbours will store into (they do so only after this rank's "set up" signal)
		for (int c = 0; c < 2; c++) {
			FS3_CUDA(cudaMemsetAsync((REAL*)s->p[c] + s->rb.plane * (size_t)s->z0, 0, s->rb.plane * sizeof(REAL), s->stream));
			FS3_CUDA(cudaMemsetAsync((REAL*)s->p[c] + s->rb.plane * (size_t)(s->z1 + 1), 0, s->rb.plane * sizeof(REAL), s->stream));
		}
		if (fs3i_sor_two_plane(s)) {   // ... and the second ghost plane of black
			if (s->rank > 0) FS3_CUDA(cudaMemsetAsync((REAL*)s->p[1] + s->rb.plane * (size_t)(s->z0 - 1), 0, s->rb.plane * sizeof(REAL), s->stream));
			if (s->rank < s->world - 1) FS3_CUDA(cudaMemsetAsync((REAL*)s->p[1] + s->rb.plane * (size_t)(s->z1 + 2), 0, s->rb.plane * sizeof(REAL), s->stream));
		}
	}
	s->system_valid = true;
	s->sor_list_valid = false;
	return FS3_OK;
}
int fs3i_project_setup(fs3_sim* s, float dt, bool from_old, bool with_gravity) {
	int rc = fs3i_alloc_pressure(s);
	if (rc) return rc;
	return s->alloc_precision ? project_setup_t<float>(s, dt, from_old, with_gravity) : project_setup_t<double>(s, dt, from_old, with_gravity);
}

template <typename REAL> static int project_gradient_t(fs3_sim* s, float dt, bool from_old, bool with_gravity) {
	Sys<REAL> sy = make_sys<REAL>(s);
	const float *sU = from_old ? s->oU : s->U, *sV = from_old ? s->oV : s->V, *sW = from_old ? s->oW : s->W;
	REAL dx = (REAL)1.0 / s->cpm;
	float gdt = s->prm.gravity_y * dt;
	const size_t plane = (size_t)s->nx * s->ny, c0 = plane * s->z0, c1 = plane * s->z1;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_PROJECT_TOVEL], s->stream);
	REAL gscale = dt / (s->prm.rho * dx);               // S3D.cpp:1015
	k_project_gradient<REAL><<<cdiv((long long)(c1 - c0), 256), 256, 0, s->stream>>>(sy, sU, sV, sW, s->phi, s->U, s->V, s->W, gscale, gdt,
	                                                                                 with_gravity ? 1 : 0, (REAL)1000.0, c0, c1);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_project_gradient(fs3_sim* s, float dt, bool from_old, bool with_gravity) {
	return s->alloc_precision ? project_gradient_t<float>(s, dt, from_old, with_gravity) : project_gradient_t<double>(s, dt, from_old, with_gravity);
}

// the whole projection on one GPU
int fs3i_project(fs3_sim* s, float dt, bool from_old, bool with_gravity) {
	int rc = fs3i_project_setup(s, dt, from_old, with_gravity);
	if (rc) return rc;
	if (s->prm.record_stage_times) cudaEventRecord(s->ev[FS3_T_PROJECT_SOR], s->stream);
	if ((rc = fs3i_sor_loop(s, s->prm.sor_iterations))) return rc;
	return fs3i_project_gradient(s, dt, from_old, with_gravity);
}

int fs3i_clear_pressure(fs3_sim* s) {
	int nb = cdiv((long long)s->rb.total, 256);
	if (s->alloc_precision) k_clear_p<float><<<nb, 256, 0, s->stream>>>(make_sys<float>(s));
	else k_clear_p<double><<<nb, 256, 0, s->stream>>>(make_sys<double>(s));
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}
int fs3i_export_pressure(fs3_sim* s, double* p, double* b, double* dg) {
	int nb = cdiv((long long)s->nC, 256);
	if (s->alloc_precision) k_export_system<float><<<nb, 256, 0, s->stream>>>(make_sys<float>(s), p, b, dg);
	else k_export_system<double><<<nb, 256, 0, s->stream>>>(make_sys<double>(s), p, b, dg);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}

// AddBodyForces as a stand-alone stage (S3D.cpp:780-788); fs3_step folds it into the projection kernels.
__global__ void k_bodyforces(float* __restrict__ V, size_t n, float gdt) {
	size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (q < n) V[q] = __fadd_rn(V[q], gdt);
}
int fs3i_bodyforces(fs3_sim* s, float dt) {
	k_bodyforces<<<cdiv((long long)s->nV, 256), 256, 0, s->stream>>>(s->V, s->nV, s->prm.gravity_y * dt);
	FS3_LAUNCH_CHECK();
	return FS3_OK;
}
