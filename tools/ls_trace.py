"""Per-item timeline of the level-set strip pipeline (debug side build with -DFS3_LS_TRACE): python tools/ls_trace.py [n=256]  (needs a GPU)"""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200 import build, sim as simmod  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
BZ = int(sys.argv[2]) if len(sys.argv) > 2 else 8
EXP = int(sys.argv[3]) if len(sys.argv) > 3 else 0
lib = build.build(extra_flags=["-DFS3_LS_TRACE", "-DLS_BZ=%d" % BZ, "-DLS_EXP=%d" % EXP], suffix="_trace%d_%d" % (BZ, EXP))
simmod.LIB_PATH = lib
s = simmod.FluidSim3(n, n, n, float(n))
s.Simulate(0.01)
s.sync()
s.advect(); s.bin(); s.levelset(); s.sync()
nty, ntz = (n + 31) // 32, (n + BZ - 1) // BZ
nt = nty * ntz
raw = np.fromfile("/tmp/ls_trace.bin", np.uint8)
tr = raw[:8 * nt * 8 * 8].view(np.uint64).reshape(-1, 8).astype(np.int64)
acc = tr[:, 4:8].copy()
tr = tr[:, :4]
ordr = raw[8 * nt * 8 * 8:].view(np.uint32)
m = (ordr >> 27).astype(int); ty = ((ordr >> 9) & 511).astype(int); tz = ((ordr >> 18) & 511).astype(int)
t0 = tr[:, 0].min()
tr = (tr - t0) / 1e3     # us
Ttot = n + 31 + BZ - 1
print("BZ", BZ, "EXP", EXP)
print("items %d, span %.1f us; steps per item %d" % (tr.shape[0], tr[:, 3].max(), Ttot))
run = tr[:, 3] - tr[:, 2]
print("step loop: mean %.2f us per item = %.1f ns per step (min %.1f, max %.1f ns)" % (run.mean(), 1e3 * run.mean() / Ttot, 1e3 * run.min() / Ttot, 1e3 * run.max() / Ttot))
print("ticket -> deps of previous sweep ready: mean %.2f us, median %.2f" % ((tr[:, 1] - tr[:, 0]).mean(), np.median(tr[:, 1] - tr[:, 0])))
print("deps ready -> first prefetch allowed:    mean %.2f us, median %.2f" % ((tr[:, 2] - tr[:, 1]).mean(), np.median(tr[:, 2] - tr[:, 1])))
for sw in range(8):
    sel = m == sw
    print("sweep %d: loop start %.1f .. %.1f us, loop end %.1f .. %.1f us" % (sw, tr[sel, 2].min(), tr[sel, 2].max(), tr[sel, 3].min(), tr[sel, 3].max()))
# hand-off lag in sweep 0 along z (tile (0, k) vs (0, k-1)) and along y
idx = {(int(a), int(b), int(c)): i for i, (a, b, c) in enumerate(zip(m, ty, tz))}
lz = [tr[idx[(0, 0, k)], 2] - tr[idx[(0, 0, k - 1)], 2] for k in range(1, ntz)]
ly = [tr[idx[(0, k, 0)], 2] - tr[idx[(0, k - 1, 0)], 2] for k in range(1, nty)]
print("sweep 0 start lag per z hop: mean %.2f us %s" % (np.mean(lz), np.round(lz[:8], 2)))
print("sweep 0 start lag per y hop: mean %.2f us %s" % (np.mean(ly), np.round(ly[:8], 2)))
