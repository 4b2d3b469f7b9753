#!/usr/bin/env python
"""Times the consumer-interface kernels and the state file at bench size on one B200.

    python tools/consumer_bench.py [--grid 256] [--reps 20] [--state /tmp/s.fs3]

Per kernel: CUDA events on the library's stream, `reps` launches after 3 warm-ups, algorithmic bytes per launch and the
fraction of the measured HBM copy peak (MEASURED_PEAKS.json).  One JSON line per kernel.
  blur_phi       8 B/cell        (phi read once, blurred copy written once; neighbour taps come from L1/L2)
  export_binned  56 B/particle + 44 B/cell   (pos+vel float4 read 32, 24 B record written; cell_start rows, counts, offsets)
  sample_phi     16 B/point + gathered taps  (12 B coordinates read, 4 B written; taps are cache hits for coherent rays)
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=256)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--points", type=int, default=1 << 24)
    ap.add_argument("--state", default=None, help="also time fs3_save_state / fs3_load_state to this path")
    args = ap.parse_args()
    import torch
    from fluidsimulation_b200 import FluidSim3
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    n = args.grid
    sim = FluidSim3(n, n, n, float(n))
    sim.Simulate(0.01)
    sim.sync()
    lib, h = sim._lib, sim._h
    stream = torch.cuda.ExternalStream(sim.stream)
    P, N = sim.num_particles, n ** 3

    def timed(fn, reps):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sim.sync()
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        sim.sync()
        return e0.elapsed_time(e1) / reps

    def report(name, ms, alg_bytes, note):
        gbs = alg_bytes / (ms * 1e-3) / 1e9
        print(json.dumps({"kernel": name, "grid": n, "ms": ms, "algorithmic_bytes": alg_bytes, "achieved_GBps": gbs,
                          "peak_GBps": peak, "frac": gbs / peak, "note": note}), flush=True)

    report("k_blur_phi_v4", timed(lambda: sim._ck(lib.fs3_blur_phi(h)), args.reps), 8.0 * N, "8 B/cell")
    a, b = C.c_void_p(), C.c_void_p()
    report("fs3_export_binned (k_export_counts + k_scan_lookback + k_export_binned)",
           timed(lambda: sim._ck(lib.fs3_export_binned(h, C.byref(a), C.byref(b))), max(3, args.reps // 4)),
           56.0 * P + 44.0 * N, "56 B/particle (32 read as 2 float4, 24 written) + 44 B/cell (4 half-rows x 8 B of cell_start, count 4+4, offset 4+4)")
    pts = torch.rand((args.points, 3), device="cuda", dtype=torch.float32)
    out = torch.empty(args.points, device="cuda", dtype=torch.float32)
    torch.cuda.synchronize()
    report("k_sample_phi (uniform random points)",
           timed(lambda: sim._ck(lib.fs3_sample_phi_device(h, C.c_void_p(pts.data_ptr()), C.c_longlong(args.points), 0, C.c_void_p(out.data_ptr()))), args.reps),
           16.0 * args.points, "16 B/point streamed; the 8 taps per point are random gathers into the %d MB phi (mostly L2 hits)" % (4 * N >> 20))
    if args.state:
        t0 = time.perf_counter(); sim.save_state(args.state); t1 = time.perf_counter(); sim.load_state(args.state); t2 = time.perf_counter()
        size = os.path.getsize(args.state)
        print(json.dumps({"op": "fs3_save_state / fs3_load_state", "grid": n, "file_bytes": size, "save_s": t1 - t0, "load_s": t2 - t1,
                          "save_GBps": size / (t1 - t0) / 1e9, "load_GBps": size / (t2 - t1) / 1e9,
                          "note": "host path: D2H copy + checksum + fwrite / fread + checksum + H2D copy"}), flush=True)
        os.remove(args.state)
    sim.close()


if __name__ == "__main__":
    main()
