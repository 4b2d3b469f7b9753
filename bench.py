#!/usr/bin/env python
"""bench.py -- BASELINE.json metric: particle-steps/s of the FluidSim3 step at 256^3 (dam break from
ResetSimulation, ~65.5M particles), SOR sweep GB/s against the HBM roofline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--grid 256] [--precision fp64|fp32]
  python bench.py --impl reference ...      times the reference's own CPU FluidSim3 (oracle/_ref) instead

A "step" is one FluidSim3::Simulate(dt) over the resident dam-break state.  `value` is timed with CUDA
events on the library's stream with the state resident in HBM; `e2e` goes through the reference-facing
calls with HOST buffers: upload m_particles from pinned memory, Simulate, read m_particles back.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

# bytes per cell update -- p 2 reads + 1 write, b, diag, 2x phi (fp32: 12+4+4+8; fp64 solve: 24+8+8+8).
# The units one launch processes are the FLUID cells of one colour (the reference skips air with `if phi<0`,
# Simulation3D.cpp:957; so does the kernel), counted from the level set outside the timed region.
SOR_BYTES_PER_CELL_ITER = {"fp64": 48.0, "fp32": 28.0}


def n_particles(n):
    return 8 * (n - 2) ** 2 * (n - 1 - n // 2)


def measured_peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 8:
                self.rows.append(f)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(int(float(r[1])) for r in self.rows if r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if r[4 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def cpu_reference_run(n, steps, warmup):
    """Times the reference CPU FluidSim3 (oracle/_ref, else the C port) -- single-threaded by construction."""
    import oracle_lib
    oracle_lib.build_oracle()
    kind = "reference" if oracle_lib.have_ref() else "port"
    sim = (oracle_lib.RefSim if kind == "reference" else oracle_lib.Oracle)(n)
    for _ in range(warmup):
        sim.step(0.01)
    t0 = time.perf_counter()
    for _ in range(steps):
        sim.step(0.01)
    dt = time.perf_counter() - t0
    P = sim.num_particles
    sim.close()
    return P * steps / dt, dt, kind, P


CPU_STEP_RATE = 1.8e6    # particle-steps/s of the reference on one host core, conservative (measured 2.2e6 - 2.6e6 on the GPU boxes)


def workload_name(n, P, total=False, extra=""):
    return "FluidSim3 dam-break full step, %d^3 grid, %d particles%s, 100 SOR iterations, dt=0.01%s" % (n, P, " in total" if total else "", extra)


def run_reference(args, rank, world):
    """The reference's own CPU FluidSim3 on the bench grid itself.  One 256^3 step takes ~30 s on one host core (the reference
    has no threads), so the run is a BOUNDED sample of the workload: as many of the K steps as fit ~2 minutes, at least one,
    from the same ResetSimulation state, no warm-up steps (nothing to warm on the CPU)."""
    if rank != 0:
        return
    n = args.grid
    est = n_particles(n) / CPU_STEP_RATE
    steps_run = int(max(1, min(args.steps, 130.0 // est)))
    value, dt, kind, P = cpu_reference_run(n, steps_run, 0)
    sample = "%d^3 dam break (%d particles), %d of the %d Simulate(0.01) steps timed (%.1f s), no warm-up, 1 thread (the reference has no threads)" % (
        n, P, steps_run, args.steps, dt)
    line = {"impl": "reference", "metric": "particle-steps/s", "value": value, "unit": "particle-steps/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps_run, "higher_is_better": True,
            "scaling": "strong" if args.gpus > 1 else "none (1 GPU)", "vs_baseline": None, "dtype": "f32 particles/grids, f64 pressure solve", "data": "synthetic",
            "config": {"workload": workload_name(n, P, args.gpus > 1), "parallelism": "reference CPU code, 1 host thread", "steps_timed": steps_run},
            "cpu_baseline": {"value": value, "unit": "particle-steps/s", "cores": 1, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "particle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


CHECK_STEPS = 2
M64 = 1 << 64


def state_check(sim, world, dist, torch):
    """fs3_checksum of this rank's share, summed over the ranks modulo 2^64 (int64 tensors wrap) -> the whole simulation's fingerprint."""
    c = sim.checksum()
    ints = [c["particle_hash"], c["cell_hash"], c["particles"]] + c["grid_hash"]
    sums = list(c["grid_sum"])
    if world > 1:
        t = torch.tensor([v - M64 if v >= (1 << 63) else v for v in ints], device="cuda", dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ints = [int(v) % M64 for v in t.tolist()]
        f = torch.tensor(sums, device="cuda", dtype=torch.float64)
        dist.all_reduce(f, op=dist.ReduceOp.SUM)
        sums = f.tolist()
    return {"particle_hash": "%016x" % ints[0], "cell_hash": "%016x" % ints[1], "particles": ints[2],
            "grid_hash": ["%016x" % v for v in ints[3:7]], "sumU": sums[0], "sumV": sums[1], "sumW": sums[2], "sumPhi": sums[3]}


def verify_sharded(sim, n, world, rank, local_rank, dist, torch, FluidSim3):
    """bench.py --gpus N verifies what it times: CHECK_STEPS steps of the sharded simulation from ResetSimulation, fingerprinted
    on the device, must equal the same steps of a single-GPU simulation (run by rank 0 on its own GPU beside its slab): the
    sharded step is bit-identical to the single-GPU step by design, so equality of the 64-bit fingerprints is the bar.
    Afterwards the sharded simulation is reset, so the timed steps start from the same state as at N = 1."""
    for _ in range(CHECK_STEPS):
        sim.Simulate(0.01)
    sim.sync()
    got = state_check(sim, world, dist, torch)
    verdict = [None]
    if rank == 0:
        if n > 384:
            verdict = [{"status": "skipped", "why": "a %d^3 single-GPU twin does not fit beside rank 0's slab" % n, "steps": CHECK_STEPS, "sharded": got}]
        else:
            one = FluidSim3(n, n, n, float(n), device=local_rank)
            for _ in range(CHECK_STEPS):
                one.Simulate(0.01)
            one.sync()
            ref = state_check(one, 1, dist, torch)
            one.close()
            keys = ("particle_hash", "cell_hash", "particles", "grid_hash")
            ok = all(got[k] == ref[k] for k in keys)
            verdict = [{"status": "pass" if ok else "FAIL", "against": "single-GPU run of the same %d steps in the same process (rank 0)" % CHECK_STEPS,
                        "steps": CHECK_STEPS, "sharded": got, "single_gpu": ref}]
    dist.broadcast_object_list(verdict, src=0)
    sim.ResetSimulation()
    sim.sync()
    return verdict[0]


def run_gpu(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from fluidsimulation_b200 import FluidSim3

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    n = args.grid
    sharded = world > 1 and not args.replicas
    if sharded:
        # z-slab sharding: ONE simulation over all ranks (strong scaling); handle blobs travel over torch.distributed once
        from fluidsimulation_b200 import torch_allgather
        sim = FluidSim3(n, n, n, float(n), device=local_rank, rank=rank, world=world, allgather=torch_allgather())
    else:
        sim = FluidSim3(n, n, n, float(n), device=local_rank)
    sim.set_params(precision=1 if args.precision == "fp32" else 0, record_stage_times=1)
    if args.gpu_formulation:
        # NOT the BASELINE.json metric: the reference's GPU formulation of the level set and the extrapolation
        sim.set_params(levelset_mode=1, extrapolation_mode=1)
    P = sim.num_particles
    stream = torch.cuda.ExternalStream(sim.stream, device=local_rank)
    check = None
    if sharded and not args.gpu_formulation:
        check = verify_sharded(sim, n, world, rank, local_rank, dist, torch, FluidSim3)
        if check["status"] == "FAIL":
            if rank == 0:
                print(json.dumps({"error": "sharded result differs from the single-GPU result", "check": check}), flush=True)
            sim.close()
            dist.destroy_process_group()
            raise SystemExit(3)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing: W warm-up + K timed steps, CUDA events on the library's stream -------
    for _ in range(args.warmup):
        sim.Simulate(0.01)
    sim.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = sim.kernel_launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sor_ms, stage_acc = 0.0, {}
    e0.record(stream)
    for _ in range(args.steps):
        sim.Simulate(0.01)
        st = sim.stage_times()      # waits for the step; event reads only, no extra device work
        sor_ms += st["PROJECT_SOR"]
        for k, v in st.items():
            stage_acc[k] = stage_acc.get(k, 0.0) + v
    e1.record(stream)
    sim.sync()
    barrier()
    clocks = sampler.stop()
    launches = sim.kernel_launches - launches0
    if world > 1:
        t = torch.tensor([float(launches)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        launches = int(t.item())
    ms = e0.elapsed_time(e1)
    final_state = state_check(sim, world if sharded else 1, dist, torch)     # outside the timed region
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    jobs = 1 if sharded else world          # sharded: P is the particle count of the whole simulation
    value = jobs * P * args.steps / (ms * 1e-3)

    # ---- end to end through the reference-facing calls with HOST buffers --------------------------------
    # m_particles lives in pinned host memory: upload it, Simulate, read it back (Old3D.cpp:234,263-269).
    e2e_steps = max(1, min(args.steps, 5))
    if sharded:
        # every rank moves the particles it owns (records + creation ids) through its own pinned buffers
        capl = int(sim.num_local_particles * 1.25) + 65536
        host = torch.empty((capl, 6), dtype=torch.float32, pin_memory=True)
        hids = torch.empty((capl,), dtype=torch.int32, pin_memory=True)
        hnp, inp = host.numpy(), hids.numpy()
        lp, li = sim.get_local_particles(out=hnp, ids=inp)
        moved = 0
        barrier()
        t0 = time.perf_counter()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        for _ in range(e2e_steps):
            sim.set_local_particles(lp, li)
            moved += lp.shape[0] * 28
            sim.Simulate(0.01)
            lp, li = sim.get_local_particles(out=hnp, ids=inp)
            moved += lp.shape[0] * 28
        g1.record(stream)
        sim.sync()
        barrier()
        t = torch.tensor([float(moved)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        bytes_each_way = float(t.item()) / 2 / e2e_steps
    else:
        host = torch.empty((P, 6), dtype=torch.float32, pin_memory=True)
        hnp = host.numpy()
        sim.get_particles(out=hnp)
        barrier()
        t0 = time.perf_counter()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        for _ in range(e2e_steps):
            sim.set_particles(hnp)
            sim.Simulate(0.01)
            sim.get_particles(out=hnp)
        g1.record(stream)
        sim.sync()
        barrier()
        bytes_each_way = P * 24
    e2e_ms = max(g0.elapsed_time(g1), (time.perf_counter() - t0) * 1e3)
    if world > 1:
        t = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    e2e_value = jobs * P * e2e_steps / (e2e_ms * 1e-3)

    # ---- roofline of the dominant kernel (one colour sweep of red-black SOR) ----------------------------
    iters = sim.get_params().sor_iterations
    n_launch = 2 * iters * args.steps
    avg_launch_ms = sor_ms / n_launch
    phi = sim.get_phi().reshape(n, n * n)
    fluid_cells = int((phi[sim.z0:sim.z1] < 0).sum())      # the cells this rank sweeps
    if world > 1:   # the slowest rank sets the time: report the largest slab
        t = torch.tensor([float(fluid_cells)], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        fluid_cells = int(t.item())
    bytes_per_launch = SOR_BYTES_PER_CELL_ITER[args.precision] * fluid_cells / 2.0
    peak, peak_src = measured_peaks()
    achieved = bytes_per_launch / (avg_launch_ms * 1e-3) / 1e9
    traffic, traffic_src = args.traffic, "command line"
    if traffic is None and world == 1 and n == 256 and args.precision == "fp64":
        # ncu --set full capture of this kernel at this configuration, committed under profiles/ (per launch, like `achieved`)
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", "sor_traffic.json")))
            traffic, traffic_src = float(t["dram_bytes_per_launch"]), t["source"]
        except Exception:
            traffic, traffic_src = None, None
    sys_mb = 2 * (n // 2 + 3) // 4 * 4 * n * (n + 2) * (4 if args.precision == "fp32" else 8) / 2 ** 20     # both colours of p
    l2_assisted = world == 1 and 48 < sys_mb <= 256
    roofline = {"bound": "hbm", "l2_assisted": l2_assisted,
                "note": ("`achieved` counts ALGORITHMIC bytes (SURVEY.md 8(d): p 24 + b 8 + diag 8 + phi/mask 8 per fluid-cell update); at this size the two "
                         "colours of p are held in the persisting L2 carve-out across the 200 launches, so DRAM sees less than that and frac can exceed 1; "
                         "`traffic` is the DRAM side") if l2_assisted else "`achieved` counts algorithmic bytes per fluid-cell update (SURVEY.md 8(d))",
                "kernel": "k_sor_half<%s>" % ("float,4" if args.precision == "fp32" else "double,2"),
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": bytes_per_launch, "avg_launch_us": avg_launch_ms * 1e3,
                "launches_timed": n_launch, "fluid_cells": fluid_cells, "grid_cells": n ** 3 // (world if sharded else 1),
                "cell_updates_per_s": fluid_cells * iters * args.steps / (sor_ms * 1e-3),
                "grid_cells_per_s": n ** 3 // (world if sharded else 1) * iters * args.steps / (sor_ms * 1e-3),
                "share_of_step": sor_ms / ms}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline:
            # ONE step of the bench workload itself where that is <= ~40 s of CPU (256^3: ~30 s), else the largest grid that is
            cn = n if n_particles(n) / CPU_STEP_RATE <= 40.0 else 256
            cv, cdt, kind, cP = cpu_reference_run(cn, 1, 0)
            cpu = {"value": cv, "unit": "particle-steps/s", "cores": 1, "kind": kind,
                   "sample": "%d^3 dam break (%d particles), 1 Simulate(0.01) step from ResetSimulation, %.1f s, single thread (the reference has no threads)" % (cn, cP, cdt)}
        if check is None:
            check = {"status": "recorded", "against": "nothing to compare at 1 GPU: these are the single-GPU fingerprints a sharded run must reproduce",
                     "steps": args.warmup + args.steps}
        line = {"metric": "particle-steps/s", "value": value, "unit": "particle-steps/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "strong" if sharded else ("weak" if world > 1 else "none (1 GPU; N > 1 shards this same simulation: strong)"), "vs_baseline": None,
                "dtype": "f32 particles/grids, %s pressure solve" % ("f32" if args.precision == "fp32" else "f64"), "data": "synthetic",
                "config": {"workload": workload_name(n, P, sharded, "; NON-DEFAULT level set + extrapolation in the reference's GPU formulation (levelset_mode=1, extrapolation_mode=1): not the headline configuration" if args.gpu_formulation else ""),
                           "parallelism": "1 GPU" if world == 1 else ("%d z-slabs, one rank per GPU, CUDA-IPC peer stores over NVLink (SOR halo fused into the sweep, particle migration, ghost layers, slab all-gathers on a copy stream); level-set sweeps and extrapolation replicated" % world if sharded else "%d independent replicas" % world),
                           "l2": "working set (%.1f GB) larger than L2; no flush" % ((P * 64 + 40.0 * n ** 3) / 1e9)},
                "e2e": {"value": e2e_value, "unit": "particle-steps/s", "h2d_bytes_per_step": int(bytes_each_way), "d2h_bytes_per_step": int(bytes_each_way),
                        "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
                "check": check, "state_after_timed_steps": final_state,
                "stage_ms_per_step": {k: v / args.steps for k, v in stage_acc.items()},
                "stage_note": "LEVELSET_SWEEP runs on a second stream beside P2G and EXTRAPOLATE: they overlap, stage times add up to more than TOTAL"}
        print(json.dumps(line), flush=True)
    sim.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--grid", type=int, default=256)
    ap.add_argument("--precision", choices=["fp64", "fp32"], default="fp64")
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--replicas", action="store_true", help="N>1: N independent simulations instead of one z-slab sharded simulation")
    ap.add_argument("--gpu-formulation", action="store_true",
                    help="levelset_mode=1 + extrapolation_mode=1 (single GPU; a different level set -- for measuring those modes, not the headline metric)")
    ap.add_argument("--traffic", type=float, default=None, help="ncu dram bytes per launch of the dominant kernel (profiles/)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_gpu(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
