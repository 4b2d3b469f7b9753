"""Line-sweep level set (fs3_params.levelset_mode = 1): the reference's GPU formulation of
ComputeLevelSet -- closest particle among 27 cells + 24 one-directional line sweeps -- as an opt-in alternative to the
CPU class's 8 Gauss-Seidel sweeps.  Parity against the reference's shaders is UNPINNED (D3D11, no graphics stack on this
box); oracle/fs3_gpu_semantics_oracle.c restates a deterministic reading.  CPU part: the restatement's own invariants
(phi is the distance to the particle it names, never below the true distance and close to it, independent of how the
particle array is ordered up to ties).  GPU part (`-m gpu`): the CUDA kernels bit-exact to the restatement, run in a
worker process."""
import ctypes as C
import os
import subprocess
import sys
import numpy as np
import pytest

from oracle_lib import Oracle, gls_levelset, gls_extrapolate
from scenes import SCENES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def cell_coords(n):
    c = np.arange(n ** 3)
    return np.stack([c % n, (c // n) % n, c // (n * n)], 1).astype(np.float32)


def dist_to(q, cells, ids):
    d = cells - q[ids]
    return (np.sqrt(d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]) - np.float32(1.0)).astype(np.float32)


@pytest.mark.parametrize("n,steps", [(16, 0), (24, 3)])
def test_phi_is_distance_to_named_particle_and_close_to_exact(n, steps):
    ora = Oracle(n)
    for _ in range(steps):
        ora.step()
    p = ora.get_particles()
    phi, closest = gls_levelset(p, n)
    q = (p[:, :3] * np.float32(n)).astype(np.float32)
    cells = cell_coords(n)
    assert closest.min() >= 0 and closest.max() < p.shape[0]
    assert np.array_equal(phi.view(np.uint32), dist_to(q, cells, closest).view(np.uint32))
    # never below the true distance; within a fraction of a cell of it (24 line sweeps are not exact, nor are the 8 3-D sweeps)
    rng = np.random.default_rng(0)
    sample = rng.integers(0, n ** 3, 300)
    q64 = q.astype(np.float64)
    exact = np.array([np.sqrt(((q64 - cells[c].astype(np.float64)) ** 2).sum(1)).min() - 1.0 for c in sample])
    err = phi[sample] - exact
    assert err.min() > -1e-5 and err.max() < 0.5
    # cells that hold particles are seeded from their own 27-cell neighbourhood: exact there
    counts = ora.get_cell_counts()
    own = np.flatnonzero(counts[sample] > 0)
    assert np.abs(err[own]).max() < 1e-5
    # more accurate than the CPU class's level set on the same particles (the reason the GPU formulation exists)
    ora.levelset()
    assert np.abs(err).max() <= np.abs(ora.get_phi()[sample] - exact).max()


@pytest.mark.parametrize("name", sorted(SCENES))
def test_edge_scenes_run(name):
    n = 16
    p = SCENES[name](n)
    phi, closest = gls_levelset(p, n)
    assert np.isfinite(phi).all() and closest.min() >= 0
    q = (p[:, :3] * np.float32(n)).astype(np.float32)
    assert np.array_equal(phi.view(np.uint32), dist_to(q, cell_coords(n), closest).view(np.uint32))


def test_mode_field_is_in_the_abi_and_the_state_file(tmp_path):
    from fluidsimulation_b200 import Fs3Params, state_file_write, state_file_info
    assert [f[0] for f in Fs3Params._fields_[-2:]] == ["levelset_mode", "extrapolation_mode"]
    txt = open(os.path.join(ROOT, "include", "fs3.h")).read()
    assert "int levelset_mode;" in txt
    n = 16
    ora = Oracle(n)
    prm = Fs3Params()
    from fluidsimulation_b200 import load_library
    assert load_library().fs3_default_params(C.byref(prm)) == 0
    assert prm.levelset_mode == 0
    prm.levelset_mode = 1
    prm.extrapolation_mode = 1
    path = str(tmp_path / "m.fs3")
    state_file_write(path, n, ora.get_particles(), *ora.get_velocity(), ora.get_phi(), params=prm)
    info = state_file_info(path)
    assert info.params.levelset_mode == 1 and info.params.extrapolation_mode == 1 and info.params.sor_iterations == 100


def test_one_layer_extrapolation_restatement():
    """FX/gpExtrapolateParticleVelocities.hlsl: hand-checked cases of the restatement."""
    # 1-D line of 5 nodes (xs=5, ys=zs=1): every node has 4 out-of-range neighbours (y and z), which count as valid zeros
    g = np.array([2.0, 0.0, 0.0, 0.0, 6.0], np.float32)
    v = np.array([1, 0, 0, 0, 1], np.uint8)
    out = gls_extrapolate(g, v, 5, 1, 1)
    assert out[0] == 2.0 and out[4] == 6.0                                  # valid nodes are never touched
    assert out[1] == np.float32(2.0) / np.float32(5.0)                      # +x invalid, -x valid (2), four OOB zeros: 2 / 5
    assert out[2] == 0.0                                                    # no valid neighbour inside: mean of four OOB zeros
    assert out[3] == np.float32(6.0) / np.float32(5.0)
    # interior node of a 3-D grid: plain mean of its valid neighbours, in the shader's order; an isolated one becomes 0
    rng = np.random.default_rng(2)
    g = rng.standard_normal(5 * 5 * 5).astype(np.float32)
    v = np.ones(125, np.uint8)
    c = 2 + 5 * (2 + 5 * 2)
    v[c] = 0
    out = gls_extrapolate(g, v, 5, 5, 5)
    s = np.float32(0)
    for o in (1, -1, 5, -5, 25, -25):
        s = np.float32(s + g[c + o])
    assert out[c] == np.float32(s / np.float32(6.0))
    v[:] = 0
    v[0] = 1
    out = gls_extrapolate(g, v, 5, 5, 5)
    assert out[c] == 0.0 and out[0] == g[0] and out[1] == np.float32(g[0] / np.float32(3.0))   # node (1,0,0): valid -x, two OOB zeros


@pytest.mark.gpu
def test_gpu_line_sweep_level_set_bitexact():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "gls_worker.py"), "16", "24", "32"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "GLS_WORKER_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-3000:]


@pytest.mark.gpu
def test_gpu_one_layer_extrapolation_bitexact():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "gex_worker.py")], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "GEX_WORKER_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-3000:]
