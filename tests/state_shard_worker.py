"""World-size-2 gloo worker (CPU only): the one-file-per-rank contract of the state format.  Every rank writes the particles
of its z-slab (with creation ids) plus the whole grids, as fs3_save_state does on a sharded handle; every rank then reads ALL
files back and must recover the original simulation state bit for bit."""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from fluidsimulation_b200 import shard_plan, state_file_read, state_file_write  # noqa: E402
from oracle_lib import Oracle  # noqa: E402


def main():
    out = sys.argv[1]
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 16
    ora = Oracle(n)
    ora.step()
    p, (u, v, w), phi = ora.get_particles(), ora.get_velocity(), ora.get_phi()
    cell = ora.get_cell_indices()                       # the reference's cell rule decides the owner (Simulation3D.cpp:286-288)
    z0, z1 = shard_plan(n, world, rank)
    cz = cell // (n * n)
    mine = np.flatnonzero((cz >= z0) & (cz < z1)).astype(np.int32)
    counts = [None] * world
    dist.all_gather_object(counts, int(mine.size))
    assert sum(counts) == p.shape[0]
    state_file_write(os.path.join(out, "rank%d.fs3" % rank), n, p[mine], u, v, w, phi, ids=mine, steps=1, rank=rank, world=world,
                     total_particles=p.shape[0], rank_particles=counts)
    dist.barrier()
    merged = np.full_like(p, np.nan)
    for r in range(world):
        info, pr, ids, ur, vr, wr, phir = state_file_read(os.path.join(out, "rank%d.fs3" % r))
        assert (info.rank, info.world, info.has_ids, info.total_particles, info.steps) == (r, world, 1, p.shape[0], 1)
        assert list(info.rank_particles)[:world] == counts and info.num_particles == counts[r]
        for a, b in ((u, ur), (v, vr), (w, wr), (phi, phir)):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        merged[ids] = pr
    assert np.array_equal(merged.view(np.uint32), p.view(np.uint32))
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("state shard worker ok", counts)


if __name__ == "__main__":
    main()
