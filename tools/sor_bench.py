#!/usr/bin/env python
"""Projection microbench (BASELINE.json config 3): red-black SOR with ghost-fluid boundaries, grids 64^3..512^3,
(i) all cells fluid (the reference's debug switch, Simulation3D.cpp:851-858) and (ii) the dam-break level set after one
step.  Prints one JSON line per case: fluid-cell updates/s, grid cells/s and algorithmic GB/s (48 B per
fluid-cell update with the fp64 solve, 28 B with fp32).   python tools/sor_bench.py [--grids 64,128,256] [--lib-suffix _x]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--grids", default="64,128,256")
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--lib-suffix", default="")
    a = ap.parse_args()
    import fluidsimulation_b200.sim as sim_mod
    if a.lib_suffix:
        sim_mod.LIB_PATH = sim_mod.LIB_PATH.replace(".so", a.lib_suffix + ".so")
    import numpy as np
    for n in [int(x) for x in a.grids.split(",")]:
        for prec in ("fp64", "fp32"):
            sim = sim_mod.FluidSim3(n, n, n, float(n))
            sim.set_params(precision=1 if prec == "fp32" else 0)
            sim.Simulate(0.01)
            sim.sync()
            for case in ("dam-break", "all-fluid"):
                if case == "all-fluid":
                    sim.set_phi_all_fluid()
                    sim.project()
                fluid = int((sim.get_phi() < 0).sum())
                sim.sor_bench(a.iters)
                ms = min(sim.sor_bench(a.iters) for _ in range(3))
                bpc = 48.0 if prec == "fp64" else 28.0
                print(json.dumps({"grid": n, "case": case, "solve": prec, "iters": a.iters, "ms": ms, "us_per_half_sweep": 1e3 * ms / (2 * a.iters),
                                  "fluid_cells": fluid, "fluid_cell_updates_per_s": fluid * a.iters / (ms * 1e-3),
                                  "grid_cells_per_s": n ** 3 * a.iters / (ms * 1e-3), "algorithmic_GBps": bpc * fluid * a.iters / (ms * 1e-3) / 1e9}), flush=True)
            sim.close()


if __name__ == "__main__":
    main()
