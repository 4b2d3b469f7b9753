"""Time stamps inside the sharded one-launch SOR (debug build with -DFS3_SOR_TRACE): where a half sweep's time goes.
   torchrun --nproc-per-node 2 tools/sor_trace.py [n=128]   (needs 2 GPUs)
Prints, for half sweeps 64..95 of the 4th projection, for the first (boundary) and last (interior) CTA, ns since the half sweep's
start of: flag seen, entries swept, fence done, (grid barrier) reached, arrived, left; and the time since the previous start."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch.distributed as dist  # noqa: E402
from fluidsimulation_b200 import build, sim as simmod  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
dist.init_process_group(backend="gloo")
if rank == 0:
    build.build(extra_flags=["-DFS3_SOR_TRACE"], suffix="_sortrace")
dist.barrier()
simmod.LIB_PATH = build.LIB.replace(".so", "_sortrace.so")
from fluidsimulation_b200 import FluidSim3, torch_allgather  # noqa: E402
s = FluidSim3(n, n, n, float(n), device=local, rank=rank, world=world, allgather=torch_allgather())
for _ in range(5):
    s.Simulate(0.01)
s.sync()
dist.barrier()
