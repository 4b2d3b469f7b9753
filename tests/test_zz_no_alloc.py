"""fs3_divergence_no_alloc (include/fs3.h): the allocation-free variant of the PrintDivergence metric, against the existing entry
point fs3_get_divergence -- needs a B200 (`-m gpu`).  Collected last on purpose.

The entry point has no counterpart in the reference.  It takes DEVICE pointers, so
the test holds U, V, W, phi and the two output doubles in torch tensors (torch is only the owner of the device memory here) and
calls through the C ABI with their addresses.  Bars: the maximum is the same fp32 expression over the same cells, so it must be
EQUAL; the L2 norm adds the same block partials in the same order, held to the 1e-12 relative bar test_gpu_parity.py uses for it."""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def no_alloc_divergence(lib, torch, n, grids, out):
    u, v, w, phi = grids
    rc = lib.fs3_divergence_no_alloc(0, None, n, n, n, C.c_void_p(u.data_ptr()), C.c_void_p(v.data_ptr()), C.c_void_p(w.data_ptr()),
                                     C.c_void_p(phi.data_ptr()), C.c_void_p(out.data_ptr()))
    assert rc == 0, lib.fs3_last_error(None).decode()
    torch.cuda.synchronize()
    return tuple(out.cpu().tolist())


def upload(torch, sim):
    return tuple(torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (*sim.get_velocity(), sim.get_phi()))


def test_no_alloc_divergence_matches_existing_entry_point_twice():
    import torch
    from fluidsimulation_b200 import FluidSim3, load_library
    lib = load_library()
    out = torch.full((2,), -1.0, dtype=torch.float64, device="cuda")
    # two calls in a row on different states and SIZES: the second, smaller call runs over the stale partial sums the first one
    # left in the shared scratch array and must not see them
    for n, steps in ((48, 2), (24, 1)):
        sim = FluidSim3(n, n, n, float(n))
        for _ in range(steps):
            sim.Simulate(0.01)
        sim.sync()
        want = sim.divergence()                     # the existing entry point (fs3_get_divergence)
        got = no_alloc_divergence(lib, torch, n, upload(torch, sim), out)
        assert want[1] > 0.0, "the test state must have a non-zero divergence"
        assert got[0] == want[0], (n, got, want)
        assert abs(got[1] - want[1]) <= 1e-12 * max(1.0, want[1]), (n, got, want)
        # and again on the same inputs: identical bits
        again = no_alloc_divergence(lib, torch, n, upload(torch, sim), out)
        assert again == got
        sim.close()


def test_no_alloc_rejects_grids_beyond_the_fixed_scratch():
    import torch
    from fluidsimulation_b200 import load_library
    lib = load_library()
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = C.c_void_p(dummy.data_ptr())
    rc = lib.fs3_divergence_no_alloc(0, None, 512, 512, 512, p, p, p, p, p)     # 2^27 cells > FS3_NO_ALLOC_MAX_CELLS = 2^24; nothing is launched
    assert rc == -1
    msg = lib.fs3_last_error(None).decode()
    assert "FS3_NO_ALLOC_MAX_CELLS" in msg and "134217728" in msg
    assert lib.fs3_divergence_no_alloc(0, None, 16, 16, 16, None, p, p, p, p) == -1
    torch.cuda.synchronize()
