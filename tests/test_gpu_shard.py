"""z-slab sharded step against the single-GPU path and the CPU oracle -- needs a B200 (`-m gpu`).

The lockstep group runs every rank of the sharded simulation inside one process on ONE device (include/fs3.h
fs3_group_*): the same slab kernels, SOR halo stores, particle migration and ghost-layer splicing as the multi-process
path, with the cross-rank waits satisfied by stream order.  Every stage of the sharded step does the same arithmetic in
the same order as the single-GPU step (the P2G gather walks the same sorted runs), so the bar is BIT-EXACT against the
single-GPU result, which test_gpu_parity.py holds against the oracle.
"""
import os
import subprocess
import sys
import numpy as np
import pytest

from oracle_lib import Oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def same(a, b):
    return np.array_equal(bits(a), bits(b))


def mk(n):
    from fluidsimulation_b200 import FluidSim3
    return FluidSim3(n, n, n, float(n))


def mkg(n, world):
    from fluidsimulation_b200 import FluidSim3Group
    return FluidSim3Group(n, world)


@pytest.mark.parametrize("n,world", [(16, 2), (16, 4), (24, 3), (32, 8)])
def test_group_reset_bitexact(n, world):
    g, one = mkg(n, world), mk(n)
    assert g.num_particles == one.num_particles
    assert sum(r.num_local_particles for r in g.ranks) == one.num_particles
    assert same(g.get_particles(), one.get_particles())
    for r in g.ranks:
        for a, b in zip(r.get_velocity(), one.get_velocity()):
            assert same(a, b)
    # ownership: every particle sits on the rank whose slab holds its cell
    cells = one.get_cell_indices()
    for r in g.ranks:
        mine = ~np.isnan(r.get_particles()[:, 0])
        z = cells[mine] // (n * n)
        assert mine.sum() == r.num_local_particles and (z >= r.z0).all() and (z < r.z1).all()
    g.close()


@pytest.mark.parametrize("n,world,steps", [(16, 2, 3), (16, 4, 2), (24, 2, 4), (24, 3, 2), (24, 8, 2), (32, 4, 6), (96, 4, 2)])
def test_group_steps_bitexact_vs_single_gpu(n, world, steps):
    g, one = mkg(n, world), mk(n)
    for k in range(steps):
        g.Simulate(0.01); one.Simulate(0.01)
        g.sync(); one.sync()
        assert same(g.get_particles(), one.get_particles()), "particles differ after step %d" % (k + 1)
        assert np.array_equal(g.get_cell_indices(), one.get_cell_indices())
        assert np.array_equal(g.get_pressure(), one.get_pressure()), "pressure differs after step %d" % (k + 1)
        for r in g.ranks:      # the grids are complete on every rank after a step
            for name, a, b in zip("UVW", r.get_velocity(), one.get_velocity()):
                assert same(a, b), "%s differs on rank %d after step %d" % (name, r.rank, k + 1)
            assert same(r.get_phi(), one.get_phi())
            assert np.array_equal(r.get_closest(), one.get_closest())
    assert sum(r.num_local_particles for r in g.ranks) == one.num_particles
    g.close()


def test_group_fp32_solve_bitexact_vs_single_gpu():
    n = 16
    g, one = mkg(n, 2), mk(n)
    g.set_params(precision=1); one.set_params(precision=1)
    for _ in range(2):
        g.Simulate(0.01); one.Simulate(0.01)
    g.sync(); one.sync()
    assert same(g.get_particles(), one.get_particles())
    g.close()


def test_group_against_oracle_and_teacher_forcing():
    """Sharded step from the oracle's state (set_particles / set_velocity / set_phi are per-rank uploads of the full arrays)."""
    n = 16
    g, ora = mkg(n, 2), Oracle(n)
    for step in range(4):
        g.set_particles(ora.get_particles()); g.set_velocity(*ora.get_velocity()); g.set_phi(ora.get_phi())
        g.Simulate(0.01); ora.step(0.01)
        g.sync()
        p, q = g.get_particles(), ora.get_particles()
        assert np.array_equal(g.get_cell_indices(), ora.get_cell_indices()), step
        assert float(np.abs(p - q).max()) <= 1e-5 * max(1.0, float(np.abs(q).max()))
        assert same(g.ranks[1].get_phi(), ora.get_phi())
    g.close()


def test_group_migration_moves_particles_between_slabs():
    """A dam break sloshes mostly in x; force z traffic with a strong +z velocity field and check ownership afterwards."""
    n = 16
    g, one = mkg(n, 2), mk(n)
    u, v, w = one.get_velocity()
    w = np.full_like(w, 4.0)            # 4 m/s * 0.01 s * 16 cells/m = 0.64 cells per step
    for s in (g, one):
        s.set_velocity(u, v, w)
    before = [r.num_local_particles for r in g.ranks]
    for _ in range(3):
        g.Simulate(0.01); one.Simulate(0.01)
    g.sync(); one.sync()
    after = [r.num_local_particles for r in g.ranks]
    assert sum(after) == sum(before) == one.num_particles
    assert same(g.get_particles(), one.get_particles())
    cells = one.get_cell_indices()
    for r in g.ranks:
        mine = ~np.isnan(r.get_particles()[:, 0])
        z = cells[mine] // (n * n)
        assert (z >= r.z0).all() and (z < r.z1).all()
    assert after != before, "the test is meant to move particles across the slab boundary"
    g.close()


def test_group_empty_slab_receives_immigrants():
    """A slab that holds no particle at first and is filled by migration later: every rank must keep the same live-buffer
    parity whatever its slab holds (the neighbours address its buffers with their own), else the immigrants land in the
    buffer the rank does not read."""
    n, world = 16, 4
    g, one = mkg(n, world), mk(n)
    rng = np.random.default_rng(11)
    P = 3000
    pos = rng.uniform(0.15, 0.85, (P, 3)).astype(np.float32)
    pos[:, 2] = rng.uniform(0.30, 0.45, P).astype(np.float32)      # n*Z in 4.8..7.2, cells z = 5..7: all in slab 1 (planes 4..7); slabs 0, 2, 3 are empty
    part = np.ascontiguousarray(np.concatenate([pos, np.zeros((P, 3), np.float32)], axis=1))
    u, v, w = one.get_velocity()
    u, v, w = np.zeros_like(u), np.zeros_like(v), np.full_like(w, 6.0)   # 6 m/s * 0.01 s * 16 cells/m: about one cell per step, upwards
    for s in (g, one):
        s.set_particles(part)
        s.set_velocity(u, v, w)
        s.set_params(gravity_y=0.0)
    assert [r.num_local_particles for r in g.ranks] == [0, P, 0, 0]
    seen = set()
    for k in range(6):
        g.Simulate(0.01); one.Simulate(0.01)
        g.sync(); one.sync()
        assert same(g.get_particles(), one.get_particles()), "particles differ after step %d" % (k + 1)
        assert np.array_equal(g.get_cell_indices(), one.get_cell_indices())
        seen.add(tuple(r.num_local_particles for r in g.ranks))
        for s in (g, one):
            s.set_velocity(u, v, w)          # keep the upward flow going (the projection takes most of it out)
    assert any(c[2] > 0 for c in seen), "the test is meant to move particles into an empty slab: %r" % (seen,)
    g.close()


def test_group_checksum_adds_up_to_single_gpu():
    """fs3_checksum: the ranks' fingerprints add up (mod 2^64) to the single-GPU values -- what bench.py --gpus N verifies."""
    n, world = 24, 3
    g, one = mkg(n, world), mk(n)
    for _ in range(2):
        g.Simulate(0.01); one.Simulate(0.01)
    g.sync(); one.sync()
    c1 = one.checksum()
    cs = [r.checksum() for r in g.ranks]
    M = 1 << 64
    assert sum(c["particles"] for c in cs) == c1["particles"] == one.num_particles
    assert sum(c["particle_hash"] for c in cs) % M == c1["particle_hash"]
    assert sum(c["cell_hash"] for c in cs) % M == c1["cell_hash"]
    for k in range(4):
        assert sum(c["grid_hash"][k] for c in cs) % M == c1["grid_hash"][k], k
        assert abs(sum(c["grid_sum"][k] for c in cs) - c1["grid_sum"][k]) <= 1e-9 * max(1.0, abs(c1["grid_sum"][k]))
    one.Simulate(0.01); one.sync()
    assert one.checksum()["particle_hash"] != c1["particle_hash"]
    g.close()


def test_group_reset_after_steps_and_hooks_refused():
    from fluidsimulation_b200 import Fs3Error
    n = 16
    g, one = mkg(n, 2), mk(n)
    g.Simulate(0.01)
    g.ResetSimulation()
    g.sync()
    assert same(g.get_particles(), one.get_particles())
    with pytest.raises(Fs3Error):
        g.ranks[0].advect()             # stage hooks are single-GPU only
    with pytest.raises(Fs3Error):
        g.ranks[0].Simulate(0.01)       # a lockstep rank cannot step alone
    g.close()


@pytest.mark.parametrize("world,n,steps,zflow", [(2, 24, 3, 0.0), (2, 24, 4, 4.0), (2, 32, 4, 4.0), (4, 12, 3, 2.0), (4, 64, 4, 4.0), (8, 96, 3, 2.0)])
def test_multiprocess_sharded_step(world, n, steps, zflow):
    """One process per GPU over CUDA IPC: the merged result must equal the single-GPU run bit for bit.  With zflow the grid
    velocity W is forced before every step, so particles migrate between slabs and every rank's SOR halo carries fluid.
    (The flow must not carry more than a slab's share of particles into the top slab: a slab's particle capacity is twice its even
    share, so 8 ranks get a gentler flow -- 4 m/s for 4 steps at 96^3 pushes 15 planes of particles into the 12-plane top slab.)
    Needs `world` B200s (`gpurun --gpus N`); skipped on a box with fewer (ranks spin on each other's flags inside kernels and
    must not share a device)."""
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
                          "--master-port", str(29533 + world), os.path.join(ROOT, "tests", "shard_worker.py"), str(n), str(steps), "--zflow", str(zflow)],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, "\n".join(l for l in out.stdout.splitlines() if l.startswith("rank 0")) + out.stdout[-1500:] + out.stderr[-1500:]
    assert "SHARD_WORKER_OK" in out.stdout


def test_multiprocess_sharded_step_fp32_solve():
    """The fp32 pressure solve (float4 vectors) of a multi-process handle: the one-launch solve reads the neighbours' planes with
    16-byte system-scope loads."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29541", os.path.join(ROOT, "tests", "shard_worker.py"), "32", "3", "--zflow", "4.0", "--fp32"],
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, "\n".join(l for l in out.stdout.splitlines() if l.startswith("rank 0")) + out.stdout[-1500:] + out.stderr[-1500:]
    assert "SHARD_WORKER_OK" in out.stdout
