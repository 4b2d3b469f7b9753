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
    "fs3_create_sharded", "fs3_shard_get_handle", "fs3_shard_connect", "fs3_shard_plan", "fs3_shard_info",
    "fs3_num_local_particles", "fs3_get_local_particles", "fs3_set_local_particles", "fs3_group_create", "fs3_group_reset", "fs3_group_step",
    "fs3_phi_device", "fs3_blur_phi", "fs3_phi_blurred_device", "fs3_get_phi_blurred", "fs3_export_binned", "fs3_get_binned",
    "fs3_sample_phi", "fs3_sample_phi_device",
    "fs3_save_state", "fs3_load_state", "fs3_step_count", "fs3_state_file_info", "fs3_state_file_write", "fs3_state_file_read",
    "fs3_state_checksum", "fs3_default_params", "fs3_get_divergence", "fs3_checksum", "fs3_divergence_no_alloc",
]
SHARD_HANDLE_BYTES = 128


class Fs3Params(C.Structure):
    _fields_ = [("sor_iterations", C.c_int), ("omega", C.c_double), ("precision", C.c_int), ("force_dt_0_01", C.c_int),
                ("record_residuals", C.c_int), ("rho", C.c_float), ("gravity_y", C.c_float), ("nu", C.c_float),
                ("particle_radius", C.c_float), ("record_stage_times", C.c_int), ("levelset_mode", C.c_int),
                ("extrapolation_mode", C.c_int)]


class Fs3StateInfo(C.Structure):
    """include/fs3.h fs3_state_info: the header of an on-disk state file."""
    _fields_ = [("version", C.c_int), ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int), ("cells_per_meter", C.c_float),
                ("num_particles", C.c_longlong), ("total_particles", C.c_longlong), ("steps", C.c_longlong),
                ("rank", C.c_int), ("world", C.c_int), ("has_ids", C.c_int), ("params", Fs3Params),
                ("rank_particles", C.c_longlong * 16), ("payload_bytes", C.c_ulonglong), ("checksum", C.c_ulonglong)]


class Fs3Check(C.Structure):
    """include/fs3.h fs3_check: order-independent fingerprints of the state (sums modulo 2^64)."""
    _fields_ = [("particle_hash", C.c_ulonglong), ("cell_hash", C.c_ulonglong), ("particles", C.c_longlong),
                ("grid_hash", C.c_ulonglong * 4), ("grid_sum", C.c_double * 4)]


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
        _lib.fs3_num_local_particles.restype = C.c_longlong
        _lib.fs3_num_local_particles.argtypes = [C.c_void_p]
        _lib.fs3_step_count.restype = C.c_longlong
        _lib.fs3_step_count.argtypes = [C.c_void_p]
        _lib.fs3_state_checksum.restype = C.c_ulonglong
        _lib.fs3_state_checksum.argtypes = [C.c_void_p, C.c_size_t, C.c_ulonglong]
        for name in ("fs3_save_state", "fs3_load_state"):
            getattr(_lib, name).argtypes = [C.c_void_p, C.c_char_p]
        _lib.fs3_default_params.argtypes = [C.POINTER(Fs3Params)]
        _lib.fs3_get_params.argtypes = [C.c_void_p, C.POINTER(Fs3Params)]
        _lib.fs3_set_params.argtypes = [C.c_void_p, C.POINTER(Fs3Params)]
        _lib.fs3_state_file_info.argtypes = [C.c_char_p, C.POINTER(Fs3StateInfo)]
        _lib.fs3_state_file_write.argtypes = [C.c_char_p, C.POINTER(Fs3StateInfo)] + [C.c_void_p] * 6
        _lib.fs3_state_file_read.argtypes = [C.c_char_p, C.POINTER(Fs3StateInfo)] + [C.c_void_p] * 6
        _lib.fs3_get_divergence.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        _lib.fs3_checksum.argtypes = [C.c_void_p, C.POINTER(Fs3Check)]
        _lib.fs3_divergence_no_alloc.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 5
        _lib.fs3_sample_phi.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]
        _lib.fs3_sample_phi_device.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_int, C.c_void_p]
    return _lib


def _host_ck(rc):
    if rc != FS3_OK:
        raise Fs3Error("fs3 error %d: %s" % (rc, load_library().fs3_last_error(None).decode()))


def state_checksum(data, seed=0):
    """64-bit checksum of a state file's payload sections (include/fs3.h fs3_state_checksum); seed chains sections."""
    a = np.ascontiguousarray(data)
    return int(load_library().fs3_state_checksum(_ptr(a), a.nbytes, seed))


def state_file_info(path):
    """Header of a state file (host only, no GPU needed)."""
    info = Fs3StateInfo()
    _host_ck(load_library().fs3_state_file_info(os.fsencode(path), C.byref(info)))
    return info


def state_file_write(path, n, particles, U, V, W, phi, ids=None, steps=0, rank=0, world=1, total_particles=None, params=None,
                     rank_particles=None):
    """Writes a state file from host arrays (host only): the same bytes fs3_save_state produces for that state."""
    p = np.ascontiguousarray(particles, np.float32).reshape(-1, 6)
    info = Fs3StateInfo()
    info.version, info.nx, info.ny, info.nz, info.cells_per_meter = 1, n, n, n, float(n)
    info.num_particles, info.steps, info.rank, info.world = p.shape[0], steps, rank, world
    info.total_particles = p.shape[0] if total_particles is None else total_particles
    info.has_ids = 0 if ids is None else 1
    if params is not None:
        info.params = params
    else:
        _host_ck(load_library().fs3_default_params(C.byref(info.params)))
    for r, c in enumerate(rank_particles if rank_particles is not None else [p.shape[0]]):
        info.rank_particles[r] = c
    ids = None if ids is None else np.ascontiguousarray(ids, np.int32)
    arrs = [np.ascontiguousarray(a, np.float32) for a in (U, V, W, phi)]
    _host_ck(load_library().fs3_state_file_write(os.fsencode(path), C.byref(info), _ptr(p), _ptr(ids), *[_ptr(a) for a in arrs]))


def state_file_read(path):
    """(info, particles, ids or None, U, V, W, phi) of a state file, checksum verified (host only)."""
    info = state_file_info(path)
    n = info.nx
    p = np.empty((info.num_particles, 6), np.float32)
    ids = np.empty(info.num_particles, np.int32) if info.has_ids else None
    U, V, W, phi = (np.empty(k, np.float32) for k in ((n + 1) * n * n, n * (n + 1) * n, n * n * (n + 1), n ** 3))
    _host_ck(load_library().fs3_state_file_read(os.fsencode(path), C.byref(info), _ptr(p), _ptr(ids), _ptr(U), _ptr(V), _ptr(W), _ptr(phi)))
    return info, p, ids, U, V, W, phi


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def shard_plan(nz, world, rank):
    """[z0, z1) of rank's slab (include/fs3.h fs3_shard_plan; pure host arithmetic, no GPU needed)."""
    lib = load_library()
    z0, z1 = C.c_int(0), C.c_int(0)
    rc = lib.fs3_shard_plan(C.c_int(nz), C.c_int(world), C.c_int(rank), C.byref(z0), C.byref(z1))
    if rc != FS3_OK:
        raise Fs3Error("z-slab sharding needs 1 <= world <= 16, nz divisible by world and slabs of at least 2 planes")
    return z0.value, z1.value


def merge_particles(parts):
    """Merges the per-rank results of get_particles() of a sharded simulation (records of other ranks are NaN)."""
    out = np.array(parts[0], copy=True)
    for p in parts[1:]:
        m = ~np.isnan(p[:, 0])
        out[m] = p[m]
    return out


def torch_allgather(group=None):
    """`allgather` callable for FluidSim3(world > 1) over torch.distributed (any backend)."""
    import torch.distributed as dist

    def f(blob):
        out = [None] * dist.get_world_size(group)
        dist.all_gather_object(out, blob, group=group)
        return out
    return f


class FluidSim3:
    """Drop-in for the reference class; grid/particle arrays cross the boundary as numpy arrays."""

    def __init__(self, xSize, ySize, zSize, CellsPerMeter, device=0, rank=0, world=1, allgather=None, _handle=None):
        """world > 1: this process is rank `rank` of a z-slab sharded simulation (one process per GPU).  `allgather` is a
        callable taking this rank's handle blob (bytes) and returning the blobs of all ranks in rank order, e.g.
        ``lambda b: torch_allgather_bytes(b)``; construction, Simulate and ResetSimulation are then collective."""
        self._lib = load_library()
        self._h = C.c_void_p()
        self.rank, self.world = rank, world
        self._owned = _handle is None
        if _handle is not None:
            self._h = _handle
        elif world == 1:
            rc = self._lib.fs3_create(C.c_int(xSize), C.c_int(ySize), C.c_int(zSize), C.c_float(CellsPerMeter),
                                      C.c_int(device), C.byref(self._h))
            if rc != FS3_OK:
                raise Fs3Error("fs3_create failed (%d): %s" % (rc, self._lib.fs3_last_error(None).decode()))
        else:
            if allgather is None:
                raise Fs3Error("a sharded FluidSim3 needs an `allgather` callable to exchange the ranks' handle blobs")
            rc = self._lib.fs3_create_sharded(C.c_int(xSize), C.c_int(ySize), C.c_int(zSize), C.c_float(CellsPerMeter),
                                              C.c_int(device), C.c_int(rank), C.c_int(world), C.byref(self._h))
            if rc != FS3_OK:
                raise Fs3Error("fs3_create_sharded failed (%d): %s" % (rc, self._lib.fs3_last_error(None).decode()))
            blob = C.create_string_buffer(SHARD_HANDLE_BYTES)
            self._ck(self._lib.fs3_shard_get_handle(self._h, blob))
            blobs = allgather(blob.raw)
            if len(blobs) != world or any(len(b) != SHARD_HANDLE_BYTES for b in blobs):
                raise Fs3Error("allgather must return one %d-byte blob per rank" % SHARD_HANDLE_BYTES)
            self._ck(self._lib.fs3_shard_connect(self._h, b"".join(blobs)))
        self.mX, self.mY, self.mZ = xSize, ySize, zSize
        n = xSize
        self.nU, self.nV, self.nW, self.nC = (n + 1) * n * n, n * (n + 1) * n, n * n * (n + 1), n ** 3
        self.z0, self.z1 = shard_plan(zSize, world, rank)

    # -- plumbing ------------------------------------------------------------------------------
    def _ck(self, rc):
        if rc != FS3_OK:
            raise Fs3Error("fs3 error %d: %s" % (rc, self._lib.fs3_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            if self._owned:
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
    def num_local_particles(self):
        return int(self._lib.fs3_num_local_particles(self._h))

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

    def get_local_particles(self, out=None, ids=None):
        """(records, creation ids) of the particles this rank owns."""
        n = self.num_local_particles
        if out is None:
            out = np.empty((n, 6), np.float32)
        if ids is None:
            ids = np.empty(n, np.int32)
        cnt = C.c_longlong(0)
        self._ck(self._lib.fs3_get_local_particles(self._h, _ptr(out), _ptr(ids), C.c_longlong(out.shape[0]), C.byref(cnt)))
        return out[:cnt.value], ids[:cnt.value]

    def set_local_particles(self, p, ids):
        p, ids = np.ascontiguousarray(p, np.float32), np.ascontiguousarray(ids, np.int32)
        self._ck(self._lib.fs3_set_local_particles(self._h, _ptr(p), _ptr(ids), C.c_longlong(p.shape[0])))

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

    def divergence(self):
        """(max, l2) of the divergence over phi < 0 cells (PrintDivergence, Simulation3D.cpp:1095-1138); sharded: this rank's planes."""
        a, b = C.c_double(0), C.c_double(0)
        self._ck(self._lib.fs3_get_divergence(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def checksum(self):
        """dict of the fs3_check fingerprints (sharded: this rank's share; shares add up modulo 2^64)."""
        c = Fs3Check()
        self._ck(self._lib.fs3_checksum(self._h, C.byref(c)))
        return {"particle_hash": int(c.particle_hash), "cell_hash": int(c.cell_hash), "particles": int(c.particles),
                "grid_hash": [int(x) for x in c.grid_hash], "grid_sum": [float(x) for x in c.grid_sum]}

    def sor_bench(self, iters):
        ms = C.c_float(0)
        self._ck(self._lib.fs3_sor_bench(self._h, iters, C.byref(ms)))
        return ms.value

    def set_phi_all_fluid(self):
        self._ck(self._lib.fs3_set_phi_all_fluid(self._h))

    # -- consumer interface (GPFluidSim's public m_gpPhiSRV / m_gpCountsSRV / m_gpBinnedParticles, Simulation.h:118-140) ----
    def blur_phi(self):
        """FX/gpBlur.hlsl into a second device buffer; returns the blurred phi as a host array."""
        self._ck(self._lib.fs3_blur_phi(self._h))
        out = np.empty(self.nC, np.float32)
        self._ck(self._lib.fs3_get_phi_blurred(self._h, _ptr(out)))
        return out

    def phi_device(self, blurred=False):
        """Raw device address of phi (or of the blurred copy) for a CUDA consumer."""
        p = C.c_void_p()
        f = self._lib.fs3_phi_blurred_device if blurred else self._lib.fs3_phi_device
        self._ck(f(self._h, C.byref(p)))
        return p.value

    def get_binned(self):
        """(records grouped by cell (P,6), offsets[n^3+1])."""
        out, off = np.empty((self.num_particles, 6), np.float32), np.empty(self.nC + 1, np.uint32)
        self._ck(self._lib.fs3_get_binned(self._h, _ptr(out), _ptr(off)))
        return out, off

    def binned_device(self):
        """Device addresses (records, offsets) of the cell-major export."""
        a, b = C.c_void_p(), C.c_void_p()
        self._ck(self._lib.fs3_export_binned(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def sample_phi(self, xyz, blurred=False):
        """FX/Render.fx map(p) at (K,3) points of the unit box: distance to the surface in metres."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.empty(xyz.shape[0], np.float32)
        self._ck(self._lib.fs3_sample_phi(self._h, _ptr(xyz), C.c_longlong(xyz.shape[0]), C.c_int(int(blurred)), _ptr(out)))
        return out

    # -- on-disk state ---------------------------------------------------------------------------------
    def save_state(self, path):
        self._ck(self._lib.fs3_save_state(self._h, os.fsencode(path)))

    def load_state(self, path):
        self._ck(self._lib.fs3_load_state(self._h, os.fsencode(path)))

    @property
    def step_count(self):
        return int(self._lib.fs3_step_count(self._h))


class FluidSim3Group:
    """All ranks of a z-slab sharded simulation inside this process on ONE device, stepped in lockstep (fs3_group_*):
    the sharded code path without needing several GPUs.  `ranks[r]` is a FluidSim3 view of rank r."""

    def __init__(self, n, world, device=0):
        self._lib = load_library()
        self.world, self.n = world, n
        self._arr = (C.c_void_p * world)()
        rc = self._lib.fs3_group_create(C.c_int(n), C.c_int(n), C.c_int(n), C.c_float(float(n)), C.c_int(device), C.c_int(world), self._arr)
        if rc != FS3_OK:
            raise Fs3Error("fs3_group_create failed (%d): %s" % (rc, self._lib.fs3_last_error(None).decode()))
        self.ranks = [FluidSim3(n, n, n, float(n), device, r, world, _handle=C.c_void_p(self._arr[r])) for r in range(world)]

    def _ck(self, rc):
        if rc != FS3_OK:
            msgs = [self._lib.fs3_last_error(C.c_void_p(self._arr[r])).decode() for r in range(self.world)]
            raise Fs3Error("fs3 group error %d: %s" % (rc, " | ".join(m for m in msgs if m)))

    def Simulate(self, dt):
        self._ck(self._lib.fs3_group_step(self._arr, C.c_int(self.world), C.c_float(dt)))

    def ResetSimulation(self):
        self._ck(self._lib.fs3_group_reset(self._arr, C.c_int(self.world)))

    def sync(self):
        for r in self.ranks:
            r.sync()

    def set_params(self, **kw):
        for r in self.ranks:
            r.set_params(**kw)

    def set_particles(self, p):
        for r in self.ranks:
            r.set_particles(p)

    def set_velocity(self, u, v, w):
        for r in self.ranks:
            r.set_velocity(u, v, w)

    def set_phi(self, phi):
        for r in self.ranks:
            r.set_phi(phi)

    def get_particles(self):
        return merge_particles([r.get_particles() for r in self.ranks])

    def get_cell_indices(self):
        out = self.ranks[0].get_cell_indices()
        for r in self.ranks[1:]:
            c = r.get_cell_indices()
            out = np.where(c >= 0, c, out)
        return out

    def get_pressure(self):
        """Natural-order p stitched from the slabs (each rank holds its own planes)."""
        n = self.n
        out = np.empty(n ** 3, np.float64)
        for r in self.ranks:
            out.reshape(n, n * n)[r.z0:r.z1] = r.get_pressure().reshape(n, n * n)[r.z0:r.z1]
        return out

    @property
    def num_particles(self):
        return self.ranks[0].num_particles

    @property
    def kernel_launches(self):
        return sum(r.kernel_launches for r in self.ranks)

    def close(self):
        for r in reversed(self.ranks):
            r._h = None
        for k in reversed(range(self.world)):
            if self._arr[k]:
                self._lib.fs3_destroy(C.c_void_p(self._arr[k]))
                self._arr[k] = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
