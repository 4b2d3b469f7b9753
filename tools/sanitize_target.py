"""Small workload for compute-sanitizer (one tool per gpurun call): a 16^3 step on one handle, a 2-rank lockstep group, the getters.
   compute-sanitizer --tool memcheck python tools/sanitize_target.py"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200 import FluidSim3, FluidSim3Group  # noqa: E402

n = 16
s = FluidSim3(n, n, n, float(n))
for _ in range(2):
    s.Simulate(0.01)
s.sync()
p = s.get_particles(); s.get_velocity(); s.get_phi(); s.get_cell_counts(); s.divergence(); s.checksum()
g = FluidSim3Group(n, 2)
for _ in range(2):
    g.Simulate(0.01)
g.sync()
q = g.get_particles()
assert (p.view("uint32") == q.view("uint32")).all(), "lockstep group differs from the single handle"
g.close(); s.close()
print("SANITIZE_TARGET_OK")
