"""Phi / binned-particle consumer interface (include/fs3.h fs3_blur_phi, fs3_export_binned,
fs3_sample_phi): the arithmetic of FX/gpBlur.hlsl, FX/gpBinParticles.hlsl and FX/Render.fx map().

Parity is UNPINNED against the reference for these three (D3D11 shaders, no graphics stack on this box, two of them racy
in the reference): the CPU part checks the C restatement (oracle/fs3_consumer_oracle.c) against independent numpy
formulations and against the pinned oracle's own cell rule; the GPU part (`-m gpu`) holds the CUDA kernels bit-exact to
the restatement.
"""
import numpy as np
import pytest

from oracle_lib import Oracle, blur_phi, sample_phi, bin_particles


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def stepped(n, steps):
    ora = Oracle(n)
    for _ in range(steps):
        ora.step()
    return ora


def sample_points(n, count, seed=0):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-0.1, 1.1, (count, 3)).astype(np.float32)         # incl. points outside the box (clamp addressing)
    k = rng.integers(0, n, (count // 4, 3))
    pts[:count // 4] = ((k + 0.5) / n).astype(np.float32)               # texel centres
    pts[count // 4] = (-5.0, 0.5, 7.0)
    return pts


# ---- CPU: the restatement against independent formulations -------------------------------------------------------
def test_blur_matches_numpy_formulation():
    n = 16
    phi = stepped(n, 1).get_phi().reshape(n, n, n)                       # [z, y, x]
    pad = np.pad(phi, 1)                                                 # out-of-range taps read 0
    c = pad[1:-1, 1:-1, 1:-1]
    acc = c + pad[1:-1, 1:-1, 2:]
    for nb in (pad[1:-1, 1:-1, :-2], pad[1:-1, 2:, 1:-1], pad[1:-1, :-2, 1:-1], pad[2:, 1:-1, 1:-1], pad[:-2, 1:-1, 1:-1]):
        acc = acc + nb
    want = (acc / np.float32(7.0)).astype(np.float32)
    assert np.array_equal(bits(blur_phi(phi.ravel(), n)), bits(want.ravel()))


def test_sample_at_texel_centres_and_linearity():
    n = 16
    phi = stepped(n, 1).get_phi()
    g = phi.reshape(n, n, n)
    # at the centre of texel k, (k + 0.5) / n, map() reads exactly texel k (t = k, weight 0)
    k = np.array([[3, 4, 5], [0, 0, 0], [15, 15, 15], [7, 0, 15]])
    pts = ((k + 0.5) / n).astype(np.float32)
    got = sample_phi(phi, n, pts)
    want = g[k[:, 2], k[:, 1], k[:, 0]] * np.float32(1.0 / n)
    assert np.allclose(got, want, rtol=1e-6, atol=1e-7)
    # the filter is linear in phi
    pts = sample_points(n, 256)
    a = sample_phi(phi, n, pts)
    b = sample_phi((2 * phi).astype(np.float32), n, pts)
    assert np.array_equal(bits(b), bits((2 * a).astype(np.float32)))
    # far outside the box the clamped border texel comes back
    far = sample_phi(phi, n, np.array([[-100.0, -100.0, -100.0], [100.0, 100.0, 100.0]], np.float32))
    assert np.allclose(far, [g[0, 0, 0] / n, g[-1, -1, -1] / n], rtol=1e-6)


def test_binned_matches_pinned_cell_rule():
    n = 16
    ora = stepped(n, 3)
    p = ora.get_particles()
    out, off = bin_particles(p, n)
    cell = ora.get_cell_indices()                                        # pinned: (int)roundf(n*X) of the reference
    counts = ora.get_cell_counts()
    assert off[0] == 0 and off[-1] == p.shape[0]
    assert np.array_equal(np.diff(off.astype(np.int64)), counts)
    # every cell holds exactly its particles (as a set of records)
    order = np.argsort(cell, kind="stable")
    want = p[order]
    key = lambda a: np.lexsort([a[:, c].view(np.uint32) for c in range(5, -1, -1)] + [np.repeat(np.arange(n ** 3), counts)])
    assert np.array_equal(bits(out[key(out)]), bits(want[key(want)]))


# ---- GPU: CUDA kernels bit-exact to the restatement ---------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("n,steps", [(16, 1), (32, 3)])
def test_gpu_blur_and_sample_bitexact(n, steps):
    from fluidsimulation_b200 import FluidSim3
    sim = FluidSim3(n, n, n, float(n))
    for _ in range(steps):
        sim.Simulate(0.01)
    phi = sim.get_phi()
    blurred = sim.blur_phi()
    assert np.array_equal(bits(blurred), bits(blur_phi(phi, n)))
    assert np.array_equal(bits(sim.get_phi()), bits(phi)), "the blur must not touch the solver's phi"
    pts = sample_points(n, 4096, seed=n)
    assert np.array_equal(bits(sim.sample_phi(pts)), bits(sample_phi(phi, n, pts)))
    assert np.array_equal(bits(sim.sample_phi(pts, blurred=True)), bits(sample_phi(blurred, n, pts)))
    assert sim.phi_device() and sim.phi_device(blurred=True) and sim.phi_device() != sim.phi_device(blurred=True)


@pytest.mark.gpu
@pytest.mark.parametrize("n,steps", [(16, 0), (24, 2), (32, 4)])
def test_gpu_binned_export_bitexact(n, steps):
    from fluidsimulation_b200 import FluidSim3
    sim = FluidSim3(n, n, n, float(n))
    for _ in range(steps):
        sim.Simulate(0.01)
    p = sim.get_particles()
    out, off = sim.get_binned()
    want, woff = bin_particles(p, n)
    assert np.array_equal(off, woff)
    assert np.array_equal(bits(out), bits(want))
    assert np.array_equal(np.diff(off.astype(np.int64)), sim.get_cell_counts())
    a, b = sim.binned_device()
    assert a and b
    assert np.array_equal(bits(sim.get_particles()), bits(p)), "the export must not disturb the particles"


@pytest.mark.gpu
@pytest.mark.parametrize("n", [16, 18])                                  # 18: the scalar (nx % 4 != 0) blur kernel
def test_gpu_blur_random_phi_bitexact(n):
    from fluidsimulation_b200 import FluidSim3
    sim = FluidSim3(n, n, n, float(n))
    phi = np.random.default_rng(n).uniform(-3, 3, n ** 3).astype(np.float32)
    sim.set_phi(phi)
    assert np.array_equal(bits(sim.blur_phi()), bits(blur_phi(phi, n)))
    pts = sample_points(n, 1024, seed=1)
    assert np.array_equal(bits(sim.sample_phi(pts)), bits(sample_phi(phi, n, pts)))


@pytest.mark.gpu
def test_gpu_consumer_at_256_properties():
    """Full bench size: size-independent properties instead of the (slow) CPU restatement."""
    from fluidsimulation_b200 import FluidSim3
    n = 256
    sim = FluidSim3(n, n, n, float(n))
    sim.Simulate(0.01)
    phi, blurred = sim.get_phi(), sim.blur_phi()
    g, bl = phi.reshape(n, n, n).astype(np.float64), blurred.reshape(n, n, n)
    pad = np.pad(g, 1)
    mean7 = (pad[1:-1, 1:-1, 1:-1] + pad[1:-1, 1:-1, 2:] + pad[1:-1, 1:-1, :-2] + pad[1:-1, 2:, 1:-1] + pad[1:-1, :-2, 1:-1]
             + pad[2:, 1:-1, 1:-1] + pad[:-2, 1:-1, 1:-1]) / 7.0
    assert float(np.abs(bl - mean7).max()) <= 1e-5 * float(np.abs(g).max())
    out, off = sim.get_binned()
    assert off[0] == 0 and off[-1] == sim.num_particles
    assert np.array_equal(np.diff(off.astype(np.int64)), sim.get_cell_counts())
    # every record sits in the cell its offset range says (cell = roundf(n*X), exact in float64)
    q = out[:, :3].astype(np.float32) * np.float32(n)
    c = np.floor(q.astype(np.float64) + 0.5).astype(np.int64)
    cell = c[:, 0] + n * (c[:, 1] + n * c[:, 2])
    assert np.array_equal(cell, np.repeat(np.arange(n ** 3), np.diff(off.astype(np.int64))))
    # and the export is a permutation of the particles: column sums agree exactly in integer arithmetic
    p = sim.get_particles()
    assert np.array_equal(bits(out).sum(0, dtype=np.uint64), bits(p).sum(0, dtype=np.uint64))


def test_restatements_match_committed_fixture():
    """tests/golden/gpuform16.npz freezes the deterministic reading of the 8(f) restatements (see make_golden_8f.py)."""
    import os
    import sys
    g = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, g)
    import make_golden_8f
    want = np.load(os.path.join(g, "gpuform16.npz"))
    got = make_golden_8f.compute()
    assert sorted(want.files) == sorted(got)
    for k in want.files:
        a, b = np.ascontiguousarray(got[k]), want[k]
        assert a.dtype == b.dtype and a.shape == b.shape, k
        assert a.tobytes() == b.tobytes(), k
