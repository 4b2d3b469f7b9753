#!/usr/bin/env python
"""Live DRAM traffic of the L2-resident SOR half sweep (k_sor_half with the persisting-L2 access window).

ncu replays kernels one at a time and does not honour the access-policy window (profiles/r01ab_*), so the DRAM bytes of the
kernel as it runs inside the solve cannot be captured with it.  This tool reads the device-wide DRAM bandwidth utilisation
from NVML's GPM counters (nvmlGpmSampleGet / NVML_GPM_METRIC_DRAM_BW_UTIL, Hopper and later) around a long back-to-back run
of the projection microbench (fs3_sor_bench: thousands of half sweeps, nothing else on the GPU), and calibrates the
"percent of theoretical peak" scale against a plain device copy of known size in the same process.

    python tools/sor_traffic_gpm.py [--grid 256] [--iters 3000] [--out profiles/sor_traffic.json]

Prints one JSON line; with --out also writes the file bench.py reads `roofline.traffic` from."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class Gpm:
    def __init__(self, index):
        import pynvml as N
        self.N = N
        N.nvmlInit()
        self.h = N.nvmlDeviceGetHandleByIndex(index)
        sup = N.nvmlGpmQueryDeviceSupport(self.h)
        if not sup.isSupportedDevice:
            raise RuntimeError("GPM not supported on this device")
        self.s = [N.nvmlGpmSampleAlloc(), N.nvmlGpmSampleAlloc()]

    def begin(self):
        self.N.nvmlGpmSampleGet(self.h, self.s[0])

    def end(self):
        N = self.N
        N.nvmlGpmSampleGet(self.h, self.s[1])
        mg = N.c_nvmlGpmMetricsGet_t()
        mg.version = N.NVML_GPM_METRICS_GET_VERSION
        mg.numMetrics = 1
        mg.sample1, mg.sample2 = self.s[0], self.s[1]
        mg.metrics[0].metricId = N.NVML_GPM_METRIC_DRAM_BW_UTIL
        N.nvmlGpmMetricsGet(mg)
        return float(mg.metrics[0].value)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grid", type=int, default=256)
    ap.add_argument("--iters", type=int, default=3000)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    from fluidsimulation_b200 import FluidSim3
    n = a.grid
    try:
        gpm = Gpm(0)
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"error": "NVML GPM unavailable: %r" % (e,)}))
        return 0
    # ---- calibration: device copy of 1 GiB -> 1 GiB, 40 times: 2 GiB of DRAM traffic per copy ------------------------------
    x = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    y = torch.empty_like(x)
    for _ in range(3):
        y.copy_(x)
    torch.cuda.synchronize()
    time.sleep(0.05)
    gpm.begin()
    t0 = time.perf_counter()
    reps = 60
    for _ in range(reps):
        y.copy_(x)
    torch.cuda.synchronize()
    t_copy = time.perf_counter() - t0
    util_copy = gpm.end()
    copy_gbs = reps * 2.0 * (1 << 30) / t_copy / 1e9
    gbs_per_percent = copy_gbs / util_copy if util_copy > 0 else float("nan")
    del x, y
    torch.cuda.empty_cache()
    # ---- the solve ---------------------------------------------------------------------------------------------------------
    sim = FluidSim3(n, n, n, float(n))
    sim.Simulate(0.01)
    sim.sync()
    fluid = int((sim.get_phi() < 0).sum())
    sim.sor_bench(200)
    time.sleep(0.05)
    gpm.begin()
    t0 = time.perf_counter()
    ms = sim.sor_bench(a.iters)              # device time of the sweep loop
    wall = time.perf_counter() - t0
    util = gpm.end()
    launches = 2 * a.iters
    us = 1e3 * ms / launches
    # utilisation is averaged over the wall interval between the two samples; the loop itself ran for `ms`
    dram_gbs = util * gbs_per_percent * (wall / (ms * 1e-3))
    bytes_per_launch = dram_gbs * 1e9 * us * 1e-6
    alg = 48.0 * fluid / 2.0
    out = {"grid": n, "kernel": "k_sor_half<double,2,0> with the persisting-L2 window, inside fs3i_sor_loop", "launches": launches,
           "us_per_launch": us, "fluid_cells": fluid, "algorithmic_bytes_per_launch": alg,
           "dram_bw_util_percent": util, "calibration": {"copy_GBps": copy_gbs, "copy_util_percent": util_copy, "GBps_per_percent": gbs_per_percent},
           "dram_GBps": dram_gbs, "dram_bytes_per_launch": bytes_per_launch, "dram_over_algorithmic": bytes_per_launch / alg,
           "source": "NVML GPM DRAM_BW_UTIL over %d back-to-back half sweeps (tools/sor_traffic_gpm.py), calibrated against a device copy; "
                     "the access-policy window is in force (ncu does not honour it)" % launches}
    print(json.dumps(out))
    if a.out:
        json.dump(out, open(a.out, "w"), indent=1)
    sim.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
