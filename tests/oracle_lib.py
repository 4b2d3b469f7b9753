"""ctypes bindings for the CPU checkers (TEST INFRASTRUCTURE ONLY).

`Oracle`  -> oracle/libfs3oracle.so  (plain-C restatement, oracle/fs3_oracle.c)
`RefSim`  -> oracle/_ref/libfs3ref.so (the unmodified reference FluidSim3, oracle/ref_capi.cpp)
Both expose the same Python surface so that tests can swap one for the other.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libfs3oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libfs3ref.so")

_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")


def build_oracle():
    """(Re)build the checkers when sources are newer than the libraries."""
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("fs3_oracle.c", "fs3_consumer_oracle.c", "fs3_gpu_semantics_oracle.c", "fs3_oracle.h")]
    if (not os.path.exists(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < max(os.path.getmtime(s) for s in srcs) \
            or (os.path.isdir("/root/reference") and not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)


def have_ref():
    return os.path.exists(REF_SO)


class _Base:
    prefix = None
    lib = None

    def _fn(self, name, restype, *argtypes):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        f.argtypes = list(argtypes)
        return f

    def __init__(self, n, cpm=None):
        self.n = n
        cpm = float(n) if cpm is None else cpm
        self.h = self._fn("create", C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float)(n, n, n, cpm)
        assert self.h, "create failed"
        self.nu, self.nv, self.nw, self.nc = (n + 1) * n * n, n * (n + 1) * n, n * n * (n + 1), n ** 3

    def close(self):
        if self.h:
            self._fn("destroy", None, C.c_void_p)(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        self._fn("reset", None, C.c_void_p)(self.h)

    def step(self, dt=0.01):
        self._fn("step", None, C.c_void_p, C.c_float)(self.h, dt)

    @property
    def num_particles(self):
        return int(self._fn("num_particles", C.c_long, C.c_void_p)(self.h))

    def get_particles(self):
        out = np.empty((self.num_particles, 6), np.float32)
        self._fn("get_particles", None, C.c_void_p, _f32p)(self.h, out)
        return out

    def set_particles(self, p):
        p = np.ascontiguousarray(p, np.float32)
        self._fn("set_particles", None, C.c_void_p, _f32p, C.c_long)(self.h, p, p.shape[0])

    def get_velocity(self):
        u, v, w = np.empty(self.nu, np.float32), np.empty(self.nv, np.float32), np.empty(self.nw, np.float32)
        self._fn("get_velocity", None, C.c_void_p, _f32p, _f32p, _f32p)(self.h, u, v, w)
        return u, v, w

    def set_velocity(self, u, v, w):
        self._fn("set_velocity", None, C.c_void_p, _f32p, _f32p, _f32p)(
            self.h, np.ascontiguousarray(u, np.float32), np.ascontiguousarray(v, np.float32),
            np.ascontiguousarray(w, np.float32))

    def get_phi(self):
        out = np.empty(self.nc, np.float32)
        self._fn("get_phi", None, C.c_void_p, _f32p)(self.h, out)
        return out

    def set_phi(self, phi):
        self._fn("set_phi", None, C.c_void_p, _f32p)(self.h, np.ascontiguousarray(phi, np.float32))

    def advect(self, dt=0.01):
        self._fn("stage_advect", None, C.c_void_p, C.c_float)(self.h, dt)

    def levelset(self):
        self._fn("stage_levelset", None, C.c_void_p)(self.h)

    def p2g(self):
        self._fn("stage_p2g", None, C.c_void_p)(self.h)

    def bodyforces(self, dt=0.01):
        self._fn("stage_bodyforces", None, C.c_void_p, C.c_float)(self.h, dt)

    def project(self, dt=0.01):
        self._fn("stage_project", None, C.c_void_p, C.c_float)(self.h, dt)

    def interpolate(self, i, j, k):
        out = np.empty(3, np.float32)
        self._fn("interpolate", None, C.c_void_p, C.c_float, C.c_float, C.c_float, _f32p)(self.h, i, j, k, out)
        return out


class RefSim(_Base):
    prefix = "ref_"

    def __init__(self, n, cpm=None):
        if RefSim.lib is None:
            RefSim.lib = C.CDLL(REF_SO)
        super().__init__(n, cpm)

    def extrapolate(self, src, valid, xs, ys, zs):
        src = np.ascontiguousarray(src, np.float32).copy()
        self._fn("stage_extrapolate", None, C.c_void_p, _f32p, _u8p, C.c_int, C.c_int, C.c_int)(
            self.h, src, np.ascontiguousarray(valid, np.uint8), xs, ys, zs)
        return src


class Oracle(_Base):
    prefix = "fs3o_"

    def __init__(self, n, cpm=None):
        if Oracle.lib is None:
            build_oracle()
            Oracle.lib = C.CDLL(ORACLE_SO)
        super().__init__(n, cpm)

    def extrapolate(self, src, valid, xs, ys, zs):
        src = np.ascontiguousarray(src, np.float32).copy()
        f = self.lib.fs3o_extrapolate
        f.restype, f.argtypes = None, [_f32p, _u8p, C.c_int, C.c_int, C.c_int]
        f(src, np.ascontiguousarray(valid, np.uint8), xs, ys, zs)
        return src

    def g2p(self, old_u, old_v, old_w, dt=0.01):
        self._fn("stage_g2p", None, C.c_void_p, C.c_float, _f32p, _f32p, _f32p)(
            self.h, dt, np.ascontiguousarray(old_u, np.float32), np.ascontiguousarray(old_v, np.float32),
            np.ascontiguousarray(old_w, np.float32))

    def project_f32(self, dt=0.01):
        self._fn("stage_project_f32", None, C.c_void_p, C.c_float)(self.h, dt)

    def get_cell_indices(self):
        out = np.empty(self.num_particles, np.int32)
        self._fn("get_cell_indices", None, C.c_void_p, _i32p)(self.h, out)
        return out

    def get_cell_counts(self):
        out = np.empty(self.nc, np.int32)
        self._fn("get_cell_counts", None, C.c_void_p, _i32p)(self.h, out)
        return out

    def get_closest(self):
        out = np.empty(self.nc, np.int32)
        self._fn("get_closest", None, C.c_void_p, _i32p)(self.h, out)
        return out

    def get_pressure(self):
        out = np.empty(self.nc, np.float64)
        self._fn("get_pressure", None, C.c_void_p, _f64p)(self.h, out)
        return out

    def get_rhs_diag(self):
        b, d = np.empty(self.nc, np.float64), np.empty(self.nc, np.float64)
        self._fn("get_rhs_diag", None, C.c_void_p, _f64p, _f64p)(self.h, b, d)
        return b, d

    def set_sor_iterations(self, it):
        self._fn("set_sor_iterations", None, C.c_void_p, C.c_int)(self.h, it)

    def set_record_residuals(self, on=True):
        self._fn("set_record_residuals", None, C.c_void_p, C.c_int)(self.h, int(on))

    def get_residuals(self):
        l2, li = np.empty(100000, np.float64), np.empty(100000, np.float64)
        n = self._fn("get_residuals", C.c_int, C.c_void_p, _f64p, _f64p, C.c_int)(self.h, l2, li, 100000)
        return l2[:n].copy(), li[:n].copy()

    def divergence(self):
        a, b = C.c_double(), C.c_double()
        self._fn("divergence", None, C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double))(
            self.h, C.byref(a), C.byref(b))
        return a.value, b.value


def _olib():
    if Oracle.lib is None:
        build_oracle()
        Oracle.lib = C.CDLL(ORACLE_SO)
    return Oracle.lib


_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")


def blur_phi(phi, n):
    """FX/gpBlur.hlsl:6-17, deterministic (Jacobi) reading -- oracle/fs3_consumer_oracle.c."""
    f = _olib().fs3o_blur_phi
    f.restype, f.argtypes = None, [_f32p, _f32p, C.c_int, C.c_int, C.c_int]
    out = np.empty(n ** 3, np.float32)
    f(np.ascontiguousarray(phi, np.float32), out, n, n, n)
    return out


def sample_phi(phi, n, xyz):
    """FX/Render.fx:77-115 map() -- oracle/fs3_consumer_oracle.c."""
    f = _olib().fs3o_sample_phi
    f.restype, f.argtypes = None, [_f32p, C.c_int, C.c_int, C.c_int, _f32p, C.c_long, _f32p]
    xyz = np.ascontiguousarray(xyz, np.float32)
    out = np.empty(xyz.shape[0], np.float32)
    f(np.ascontiguousarray(phi, np.float32), n, n, n, xyz, xyz.shape[0], out)
    return out


def bin_particles(particles, n):
    """Cell-major particle records + offsets[n^3+1] -- oracle/fs3_consumer_oracle.c."""
    f = _olib().fs3o_bin_particles
    f.restype, f.argtypes = None, [_f32p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_float, _f32p, _u32p]
    p = np.ascontiguousarray(particles, np.float32)
    out, off = np.zeros_like(p), np.empty(n ** 3 + 1, np.uint32)
    f(p, p.shape[0], n, n, n, float(n), out, off)
    return out, off


def gls_levelset(particles, n, radius=1.0):
    """Line-sweep level set (27-cell seed + 24 line sweeps) -- oracle/fs3_gpu_semantics_oracle.c.  Returns (phi, closest)."""
    f = _olib().fs3o_gls_levelset
    f.restype, f.argtypes = C.c_int, [_f32p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, _f32p, _i32p]
    p = np.ascontiguousarray(particles, np.float32)
    phi, closest = np.empty(n ** 3, np.float32), np.empty(n ** 3, np.int32)
    rc = f(p, p.shape[0], n, n, n, float(n), float(radius), phi, closest)
    assert rc == 0, "no particle inside the grid"
    return phi, closest


def gls_extrapolate(src, valid, xs, ys, zs):
    """One-layer extrapolation (FX/gpExtrapolateParticleVelocities.hlsl) -- oracle/fs3_gpu_semantics_oracle.c."""
    f = _olib().fs3o_gls_extrapolate
    f.restype, f.argtypes = None, [_f32p, _u8p, C.c_int, C.c_int, C.c_int]
    g = np.ascontiguousarray(src, np.float32).copy()
    f(g, np.ascontiguousarray(valid, np.uint8), xs, ys, zs)
    return g


def cellhash(particles, n):
    """word-wise FNV-1a over per-particle cell indices (roundf semantics)."""
    p = particles[:, :3].astype(np.float32) * np.float32(n)
    # roundf = half away from zero, computed exactly in float64 (n*p is an exact fp32 value)
    pd = p.astype(np.float64)
    c = np.where(pd >= 0, np.floor(pd + 0.5), np.ceil(pd - 0.5)).astype(np.int64)
    key = (c[:, 0] + n * (c[:, 1] + n * c[:, 2])).astype(np.uint32)
    h = 1469598103934665603
    for k in key.tolist():
        h = ((h ^ k) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h
