#!/usr/bin/env python
"""Level-set stage alone (seed + the 8 sweeps on the main stream, nothing beside them): python tools/ls_bench.py [n=256] [reps=5]"""
import os
import sys
import time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200 import FluidSim3  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
s = FluidSim3(n, n, n, float(n))
s.set_params(record_stage_times=1)
for k in range(2):
    s.Simulate(0.01)
s.sync()
print("in-step stage times (ms):", {k: round(v, 3) for k, v in s.stage_times().items()})
s.advect(); s.bin(); s.sync()
ts = []
for _ in range(reps):
    t0 = time.perf_counter()
    s.levelset()
    s.sync()
    ts.append((time.perf_counter() - t0) * 1e3)
print("levelset stage alone, %d^3: min %.3f ms, all %s" % (n, min(ts), [round(t, 3) for t in ts]))
