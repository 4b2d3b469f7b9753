"""Edge-case inputs (tests/scenes.py): collisions (hundreds of particles and exact duplicates in one cell), boundaries
(particles exactly on cell centres, half-cell faces and the clamp box), empties (300 particles in the whole box) and the
minimum particle count.  The reference has no tests of its own, so these are pinned the only way
there is: CPU part -- the C restatement against the unmodified reference (oracle/_ref) on every scene, bit for bit;
GPU part (`-m gpu`) -- the CUDA path against the restatement, stage by stage and free-running, through the C ABI."""
import numpy as np
import pytest

from oracle_lib import Oracle, RefSim, have_ref
from scenes import SCENES, grids

N = 16
RTOL = 1e-5   # north_star tolerance for the stage that sums in a different order (P2G)


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def same(a, b):
    return np.array_equal(bits(a), bits(b))


def close(a, b, rtol=RTOL):
    scale = max(1.0, float(np.abs(b).max()))
    return float(np.abs(a.astype(np.float64) - b.astype(np.float64)).max()) <= rtol * scale


def load(sim, name):
    sim.set_particles(SCENES[name](N))
    sim.set_velocity(*grids(N))


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (reference sources absent)")
@pytest.mark.parametrize("name", sorted(SCENES))
def test_oracle_matches_reference_on_edge_scenes(name):
    r, o = RefSim(N), Oracle(N)
    load(r, name); load(o, name)
    for step in range(3):
        r.step(); o.step()
        assert same(r.get_particles(), o.get_particles()), step
        for a, b in zip(r.get_velocity(), o.get_velocity()):
            assert same(a, b), step
        assert same(r.get_phi(), o.get_phi()), step
    assert np.isfinite(o.get_particles()).all()


def test_scenes_are_what_they_claim():
    o = Oracle(N)
    load(o, "crowded")
    assert o.get_cell_counts().max() > 300                      # collisions
    load(o, "sparse")
    assert (o.get_cell_counts() == 0).mean() > 0.9              # empties
    load(o, "lattice")
    q = SCENES["lattice"](N)[:, :3] * np.float32(N)
    assert (q == np.round(q)).all(axis=1).sum() >= 288          # exactly on cell centres
    assert ((q - np.floor(q)) == 0.5).any(axis=1).sum() >= 96  # exactly on half-cell faces
    assert SCENES["minimal"](N).shape[0] == 273


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SCENES))
def test_gpu_stages_on_edge_scenes(name):
    from fluidsimulation_b200 import FluidSim3
    sim, ora = FluidSim3(N, N, N, float(N)), Oracle(N)
    load(sim, name); load(ora, name)
    sim.set_phi(ora.get_phi())
    sim.advect(); ora.advect()
    assert same(sim.get_particles(), ora.get_particles()), "advect not bit-exact"
    sim.bin()
    assert np.array_equal(sim.get_cell_indices(), ora.get_cell_indices())
    assert np.array_equal(sim.get_cell_counts(), ora.get_cell_counts())
    assert same(sim.get_particles(), ora.get_particles()), "sorting must not change creation-order readback"
    sim.levelset(); ora.levelset()
    assert same(sim.get_phi(), ora.get_phi()), "phi not bit-exact"
    assert np.array_equal(sim.get_closest(), ora.get_closest()), "closest particle differs"
    sim.p2g(); ora.p2g()
    for nm, a, b in zip("UVW", sim.get_velocity(), ora.get_velocity()):
        assert close(a, b), "P2G %s: %g" % (nm, np.abs(a - b).max())
    old = ora.get_velocity()
    sim.set_velocity(*old)
    sim.bodyforces(); ora.bodyforces()
    sim.project(); ora.project()
    pg, po = sim.get_pressure(), ora.get_pressure()
    assert np.abs(pg - po).max() <= 1e-12 * max(1.0, np.abs(po).max()), "pressure: %g" % np.abs(pg - po).max()
    for nm, a, b in zip("UVW", sim.get_velocity(), ora.get_velocity()):
        assert close(a, b, 1e-7), "projected %s: %g" % (nm, np.abs(a - b).max())
    sim.set_velocity(*ora.get_velocity())
    sim.g2p(*old); ora.g2p(*old)
    assert same(sim.get_particles(), ora.get_particles()), "G2P not bit-exact"


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SCENES))
def test_gpu_free_running_on_edge_scenes(name):
    from fluidsimulation_b200 import FluidSim3
    sim, ora = FluidSim3(N, N, N, float(N)), Oracle(N)
    load(sim, name); load(ora, name)
    sim.Simulate(0.01); ora.step()
    sim.sync()
    assert np.array_equal(sim.get_cell_indices(), ora.get_cell_indices())
    assert same(sim.get_phi(), ora.get_phi())
    sim.Simulate(0.01); ora.step()
    sim.sync()
    p, q = sim.get_particles(), ora.get_particles()
    assert np.isfinite(p).all()
    assert close(p, q), "particles after 2 steps: %g" % np.abs(p - q).max()
    for a, b in zip(sim.get_velocity(), ora.get_velocity()):
        assert close(a, b)
