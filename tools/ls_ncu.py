"""Runs the level-set stage a few times (for ncu captures of k_ls_sweep_strips): python tools/ls_ncu.py [n=128]"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200 import FluidSim3  # noqa: E402
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
s = FluidSim3(n, n, n, float(n))
s.Simulate(0.01); s.sync()
s.advect(); s.bin()
for _ in range(3):
    s.levelset()
s.sync()
print("ok")
