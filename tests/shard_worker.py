"""Worker of the multi-process sharded tests: `torchrun --nproc-per-node N tests/shard_worker.py 24 3`.
Every rank builds its slab of the simulation on its own GPU, steps it, and rank 0 compares the merged result bit for
bit with a single-GPU run of the same library on its device."""
import argparse
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from fluidsimulation_b200 import FluidSim3, merge_particles, torch_allgather
    ap = argparse.ArgumentParser()
    ap.add_argument("n", type=int, nargs="?", default=24)
    ap.add_argument("steps", type=int, nargs="?", default=3)
    ap.add_argument("--zflow", type=float, default=0.0, help="m/s: W is overwritten with this value before every step, so that particles "
                    "cross slab boundaries (migration) and both SOR neighbours of every interior rank carry fluid")
    ap.add_argument("--fp32", action="store_true", help="pressure solve in fp32 (set_params precision=1) on the sharded run and on the twin")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    dist.init_process_group(backend="gloo")
    n = a.n
    sim = FluidSim3(n, n, n, float(n), device=local, rank=rank, world=world, allgather=torch_allgather())
    owned0 = sim.num_local_particles

    def run(s):
        if a.fp32:
            s.set_params(precision=1)
        for _ in range(a.steps):
            if a.zflow:
                u, v, w = s.get_velocity()
                s.set_velocity(u, v, np.full_like(w, a.zflow))
            s.Simulate(0.01)
        s.sync()
    run(sim)
    moved = [None] * world
    dist.all_gather_object(moved, sim.num_local_particles - owned0)
    mine = sim.get_particles()
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    press = [None] * world
    dist.all_gather_object(press, sim.get_pressure().reshape(n, n * n)[sim.z0:sim.z1])
    uvw = sim.get_velocity()
    ok = True
    if rank == 0:
        one = FluidSim3(n, n, n, float(n), device=local)
        run(one)
        merged = merge_particles(parts)
        same = lambda x, y: np.array_equal(np.ascontiguousarray(x).view(np.uint32), np.ascontiguousarray(y).view(np.uint32))
        checks = {"particles": same(merged, one.get_particles()), "pressure": np.array_equal(np.concatenate(press).reshape(-1), one.get_pressure())}
        for nm, x, y in zip("UVW", uvw, one.get_velocity()):
            checks[nm] = same(x, y)
        checks["phi"] = same(sim.get_phi(), one.get_phi())
        ok = all(checks.values())
        if not ok:
            print("rank 0: fields that differ from the single-GPU run:", [k for k, v in checks.items() if not v])
        print("rank 0: merged result %s the single-GPU run (%d^3, %d particles, %d steps, world %d, zflow %g, change of the ranks' particle counts %r)" % (
            "matches" if ok else "DIFFERS FROM", n, merged.shape[0], a.steps, world, a.zflow, moved))
        if a.zflow and not any(moved):
            ok = False
            print("rank 0: --zflow was meant to move particles between slabs and did not")
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    # every rank holds the complete grids
    ref = [None]
    if rank == 0:
        ref = [[x.tobytes() for x in uvw]]
    dist.broadcast_object_list(ref, src=0)
    grids_equal = all(x.tobytes() == y for x, y in zip(uvw, ref[0]))
    allok = [None] * world
    dist.all_gather_object(allok, grids_equal)
    sim.close()
    dist.barrier()
    if rank == 0 and flag[0] and all(allok):
        print("SHARD_WORKER_OK")
    dist.destroy_process_group()
    sys.exit(0 if (flag[0] and all(allok)) else 1)


if __name__ == "__main__":
    main()
