"""Per-item timeline of the level-set dataflow sweep (debug build with -DFS3_LS_TRACE).
   python tools/ls_trace.py [n=256]   (needs a GPU)"""
import ctypes as C
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fluidsimulation_b200 import build, sim as simmod  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
lib = build.build(extra_flags=["-DFS3_LS_TRACE"], suffix="_trace")
simmod.LIB_PATH = lib
out = "/tmp/ls_trace.bin"
s = simmod.FluidSim3(n, n, n, float(n))
s.Simulate(0.01)
s.sync()
os.environ["FS3_LS_TRACE_FILE"] = out
s.Simulate(0.01)
s.sync()
nt = ((n + 15) // 16) ** 2 * ((n + 7) // 8)   # LS_TX x LS_TY x LS_TZ = 16 x 16 x 8 tiles
raw = np.fromfile(out, np.uint8)
tr = raw[:8 * nt * 8 * 8].view(np.uint64).reshape(-1, 8).astype(np.int64)
quiet = tr[:, 6].copy()
ordr = raw[8 * nt * 8 * 8:].view(np.uint32)
sweep = (ordr >> 27).astype(int)
print('quiet fraction per sweep:', [round(float(quiet[sweep == m].mean()), 3) for m in range(8)], 'overall', round(float(quiet.mean()), 3))
t0 = tr[:, 0].min()
tr = tr[:, :6] - t0
names = ["ticket->self ready", "self->deps ready", "deps->loaded", "compute", "fence", ]
d = np.diff(tr, axis=1) / 1e3
print("items %d, span %.3f ms" % (tr.shape[0], (tr[:, 5].max()) / 1e6))
for k, nm in enumerate(names):
    print("%-22s mean %8.2f us  median %8.2f  p90 %8.2f  max %8.2f" % (nm, d[:, k].mean(), np.median(d[:, k]), np.percentile(d[:, k], 90), d[:, k].max()))
busy = (tr[:, 5] - tr[:, 2]) / 1e3
print("busy (deps ready -> published): mean %.2f us; sum/CTAs = %.3f ms" % (busy.mean(), busy.sum() / 1e3 / (148 * 3)))
# progress: completion time by ticket decile
q = np.linspace(0, tr.shape[0] - 1, 11).astype(int)
print("publish time (ms) at ticket deciles:", np.round(tr[q, 5] / 1e6, 3))
for nm, sel in (("quiet", quiet == 1), ("active", quiet == 0)):
    if sel.any():
        print("%-7s items %7d: load %.2f us, compute %.2f us, fence %.2f us, deps-ready -> published %.2f us" % (
            nm, int(sel.sum()), d[sel, 2].mean(), d[sel, 3].mean(), d[sel, 4].mean(), busy[sel].mean()))
print("active fraction by ticket decile:", [round(float(1 - quiet[q[i]:q[i + 1] + 1].mean()), 3) for i in range(10)])
