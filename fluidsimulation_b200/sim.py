"""Host-side mirror of the reference `FluidSim3` class over the C ABI (include/fs3.h).

Reference surface (FluidSimulationDemo/Simulation3D.h:18-24,160-162):
    FluidSim3(xSize, ySize, zSize, CellsPerMeter); ResetSimulation(); Simulate(dt); m_particles
The same names are kept here; everything runs in libfluidsim3_b200.so on the GPU.  There is no CPU
fallback: importing works without a GPU (so that symbol checks can run), constructing a simulation
without the library or without an sm_100 device raises.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfluidsim3_b200.so")

FS3_OK = 0
STAGE_NAMES = ["ADVECT", "BIN", "LEVELSET_SEED", "LEVELSET_SWEEP", "P2G", "EXTRAPOLATE", "PROJECT_SETUP", "PROJECT_SOR",
               "PROJECT_TOVEL", "FLIP_APPLY", "TOTAL"]

# every symbol include/fs3.h declares (checked against the header by tests/test_abi.py)
ABI_SYMBOLS = [
    "fs3_create", "fs3_destroy", "fs3_reset", "fs3_step", "fs3_sync", "fs3_last_error", "fs3_num_particles",
    "fs3_get_particles", "fs3_set_particles", "fs3_get_velocity", "fs3_set_velocity", "fs3_get_phi", "fs3_set_phi",
    "fs3_get_params", "fs3_set_params", "fs3_stage_advect", "fs3_stage_bin", "fs3_stage_levelset", "fs3_stage_p2g",
    "fs3_stage_bodyforces", "fs3_stage_project", "fs3_stage_g2p", "fs3_extrapolate", "fs3_get_cell_indices",
    "fs3_get_cell_counts", "fs3_get_closest", "fs3_get_pressure", "fs3_get_rhs_diag", "fs3_get_sor_residuals",
    "fs3_stage_times", "fs3_sor_bench", "fs3_set_phi_all_fluid", "fs3_stream", "fs3_kernel_launches",
]


class Fs3Params(C.Structure):
    _fields_ = [("sor_iterations", C.c_int), ("omega", C.c_double), ("precision", C.c_int), ("force_dt_0_01", C.c_int),
                ("record_residuals", C.c_int), ("rho", C.c_float), ("gravity_y", C.c_float), ("nu", C.c_float),
                ("particle_radius", C.c_float), ("record_stage_times", C.c_int)]


class Fs3Error(RuntimeError):
    pass


_lib = None


def load_library():
    """Loads the CUDA library; raises (loudly) when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Fs3Error("libfluidsim3_b200.so is missing: run `python fluidsimulation_b200/build.py` "
                           "(there is no CPU fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.fs3_last_error.restype = C.c_char_p
        _lib.fs3_last_error.argtypes = [C.c_void_p]
        _lib.fs3_num_particles.restype = C.c_longlong
        _lib.fs3_num_particles.argtypes = [C.c_void_p]
        _lib.fs3_kernel_launches.restype = C.c_longlong
        _lib.fs3_kernel_launches.argtypes = [C.c_void_p]
        _lib.fs3_stream.restype = C.c_void_p
        _lib.fs3_stream.argtypes = [C.c_void_p]
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class FluidSim3:
    """Drop-in for the reference class; grid/particle arrays cross the boundary as numpy arrays."""

    def __init__(self, xSize, ySize, zSize, CellsPerMeter, device=0):
        self._lib = load_library()
        self._h = C.c_void_p()
        rc = self._lib.fs3_create(C.c_int(xSize), C.c_int(ySize), C.c_int(zSize), C.c_float(CellsPerMeter),
                                  C.c_int(device), C.byref(self._h))
        if rc != FS3_OK:
            raise Fs3Error("fs3_create failed (%d): %s" % (rc, self._lib.fs3_last_error(None).decode()))
        self.mX, self.mY, self.mZ = xSize, ySize, zSize
        n = xSize
        self.nU, self.nV, self.nW, self.nC = (n + 1) * n * n, n * (n + 1) * n, n * n * (n + 1), n ** 3

    # -- plumbing ------------------------------------------------------------------------------
    def _ck(self, rc):
        if rc != FS3_OK:
            raise Fs3Error("fs3 error %d: %s" % (rc, self._lib.fs3_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.fs3_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- reference surface -----------------------------------------------------------------------
    def ResetSimulation(self):
        self._ck(self._lib.fs3_reset(self._h))

    def Simulate(self, dt):
        self._ck(self._lib.fs3_step(self._h, C.c_float(dt)))

    @property
    def m_particles(self):
        """(P,6) float32 rows X,Y,Z,uX,uY,uZ in creation order (ParticleDefs.h:39-47)."""
        return self.get_particles()

    # -- state access ----------------------------------------------------------------------------
    def sync(self):
        self._ck(self._lib.fs3_sync(self._h))

    @property
    def num_particles(self):
        return int(self._lib.fs3_num_particles(self._h))

    @property
    def kernel_launches(self):
        return int(self._lib.fs3_kernel_launches(self._h))

    @property
    def stream(self):
        return self._lib.fs3_stream(self._h)

    def get_particles(self, out=None):
        if out is None:
            out = np.empty((self.num_particles, 6), np.float32)
        self._ck(self._lib.fs3_get_particles(self._h, _ptr(out)))
        return out

    def set_particles(self, p):
        p = np.ascontiguousarray(p, np.float32)
        self._ck(self._lib.fs3_set_particles(self._h, _ptr(p), C.c_longlong(p.shape[0])))

    def get_velocity(self):
        u, v, w = np.empty(self.nU, np.float32), np.empty(self.nV, np.float32), np.empty(self.nW, np.float32)
        self._ck(self._lib.fs3_get_velocity(self._h, _ptr(u), _ptr(v), _ptr(w)))
        return u, v, w

    def set_velocity(self, u, v, w):
        u, v, w = (np.ascontiguousarray(a, np.float32) for a in (u, v, w))
        self._ck(self._lib.fs3_set_velocity(self._h, _ptr(u), _ptr(v), _ptr(w)))

    def get_phi(self):
        out = np.empty(self.nC, np.float32)
        self._ck(self._lib.fs3_get_phi(self._h, _ptr(out)))
        return out

    def set_phi(self, phi):
        phi = np.ascontiguousarray(phi, np.float32)
        self._ck(self._lib.fs3_set_phi(self._h, _ptr(phi)))

    def get_params(self):
        p = Fs3Params()
        self._ck(self._lib.fs3_get_params(self._h, C.byref(p)))
        return p

    def set_params(self, **kw):
        p = self.get_params()
        for k, v in kw.items():
            setattr(p, k, v)
        self._ck(self._lib.fs3_set_params(self._h, C.byref(p)))

    # -- stages ------------------------------------------------------------------------------------
    def advect(self, dt=0.01):
        self._ck(self._lib.fs3_stage_advect(self._h, C.c_float(dt)))

    def bin(self):
        self._ck(self._lib.fs3_stage_bin(self._h))

    def levelset(self):
        self._ck(self._lib.fs3_stage_levelset(self._h))

    def p2g(self):
        self._ck(self._lib.fs3_stage_p2g(self._h))

    def bodyforces(self, dt=0.01):
        self._ck(self._lib.fs3_stage_bodyforces(self._h, C.c_float(dt)))

    def project(self, dt=0.01):
        self._ck(self._lib.fs3_stage_project(self._h, C.c_float(dt)))

    def g2p(self, old_u, old_v, old_w, dt=0.01):
        a, b, c = (np.ascontiguousarray(x, np.float32) for x in (old_u, old_v, old_w))
        self._ck(self._lib.fs3_stage_g2p(self._h, C.c_float(dt), _ptr(a), _ptr(b), _ptr(c)))

    def extrapolate(self, src, valid, xs, ys, zs):
        g = np.ascontiguousarray(src, np.float32).copy()
        v = np.ascontiguousarray(valid, np.uint8)
        self._ck(self._lib.fs3_extrapolate(self._h, _ptr(g), _ptr(v), xs, ys, zs))
        return g

    # -- hooks ---------------------------------------------------------------------------------------
    def get_cell_indices(self):
        out = np.empty(self.num_particles, np.int32)
        self._ck(self._lib.fs3_get_cell_indices(self._h, _ptr(out)))
        return out

    def get_cell_counts(self):
        out = np.empty(self.nC, np.int32)
        self._ck(self._lib.fs3_get_cell_counts(self._h, _ptr(out)))
        return out

    def get_closest(self):
        out = np.empty(self.nC, np.int32)
        self._ck(self._lib.fs3_get_closest(self._h, _ptr(out)))
        return out

    def get_pressure(self):
        out = np.empty(self.nC, np.float64)
        self._ck(self._lib.fs3_get_pressure(self._h, _ptr(out)))
        return out

    def get_rhs_diag(self):
        b, d = np.empty(self.nC, np.float64), np.empty(self.nC, np.float64)
        self._ck(self._lib.fs3_get_rhs_diag(self._h, _ptr(b), _ptr(d)))
        return b, d

    def get_residuals(self):
        n = C.c_int(0)
        l2, li = np.empty(4096, np.float64), np.empty(4096, np.float64)
        self._ck(self._lib.fs3_get_sor_residuals(self._h, _ptr(l2), _ptr(li), 4096, C.byref(n)))
        return l2[:n.value].copy(), li[:n.value].copy()

    def stage_times(self):
        ms = np.zeros(len(STAGE_NAMES), np.float32)
        self._ck(self._lib.fs3_stage_times(self._h, _ptr(ms), len(STAGE_NAMES)))
        return dict(zip(STAGE_NAMES, ms.tolist()))

    def sor_bench(self, iters):
        ms = C.c_float(0)
        self._ck(self._lib.fs3_sor_bench(self._h, iters, C.byref(ms)))
        return ms.value

    def set_phi_all_fluid(self):
        self._ck(self._lib.fs3_set_phi_all_fluid(self._h))
