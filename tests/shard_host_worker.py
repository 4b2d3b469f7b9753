"""World-size-2 gloo worker (CPU only): the host-side logic of the z-slab sharding -- slab plan, handle-blob all-gather in
rank order, ownership by (int)roundf(n*Z) (Simulation3D.cpp:288) and the merge of per-rank particle read-backs."""
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200.sim import merge_particles, shard_plan, torch_allgather  # noqa: E402


def main():
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 32
    z0, z1 = shard_plan(n, world, rank)
    assert (z0, z1) == (rank * n // world, (rank + 1) * n // world)
    # the transport bench.py uses for the CUDA-IPC blobs: 128 opaque bytes per rank, delivered in rank order
    blob = bytes([rank]) * 128
    blobs = torch_allgather()(blob)
    assert len(blobs) == world and all(blobs[r] == bytes([r]) * 128 for r in range(world))
    # ownership: the same synthetic particle set on every rank, each keeps its slab, NaN elsewhere (fs3_get_particles contract)
    rng = np.random.default_rng(7)
    p = rng.uniform(-0.4 / n, 1 - 0.6 / n, size=(5000, 6)).astype(np.float32)
    q = (p[:, 2].astype(np.float32) * np.float32(n)).astype(np.float64)
    cz = np.clip(np.where(q >= 0, np.floor(q + 0.5), -np.floor(-q + 0.5)).astype(int), 0, n - 1)   # roundf: half away from zero
    mine = (cz >= z0) & (cz < z1)
    part = np.full_like(p, np.nan)
    part[mine] = p[mine]
    parts = [None] * world
    dist.all_gather_object(parts, part)
    counts = [int((~np.isnan(x[:, 0])).sum()) for x in parts]
    assert sum(counts) == p.shape[0], counts           # every particle has exactly one owner
    assert np.array_equal(merge_particles(parts), p)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("shard host worker ok", counts)


if __name__ == "__main__":
    main()
