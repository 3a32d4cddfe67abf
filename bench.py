#!/usr/bin/env python
"""bench.py -- LOWESS fitted points/sec on B200 (BASELINE.json metric).

A step = one complete smooth() (windows + nsteps+1 fit passes + nsteps
robustness updates) of the workload.  Default workload = BASELINE.json
configs[1]: single series n = 1e6, f = 0.05, 3 robustness iterations -- the
largest configuration that finishes a step in well under a minute on one GPU
(the headline n = 1e8 needs >= 15 min per step at 100 % of the FP64 roofline;
run it with --workload c4).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun ... bench.py --gpus N ...        (one rank per GPU, N > 1)

value  : whole-job points/sec, data resident in HBM, CUDA-event timed, max over ranks
e2e    : the same through the host-buffer C ABI cf_lowess_smooth (pinned host
         arrays in, host arrays out; copies inside the timed region)
roofline: the fit kernel against the MEASURED FP64 FMA peak of this GPU
cpu_baseline / --impl reference: the oracle (C restatement of the reference; the
         Fortran reference itself cannot be compiled in this image) on the host cores
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (n, f, nsteps, description)
    "c2": (1_000_000, 0.05, 3, "single series n=1e6 f=0.05 nsteps=3 (BASELINE configs[1])"),
    "c2ref": (1_000_000, 0.05, 2, "single series n=1e6 f=0.05 nsteps=2 (reference smooth() default)"),
    "c5": (2_000_000, 0.8, 5, "wide span n=2e6 f=0.8 nsteps=5 (BASELINE configs[4])"),
    "c4": (100_000_000, 0.05, 3, "single series n=1e8 f=0.05 nsteps=3 (BASELINE headline; >15 min/step)"),
    "small": (100_000, 0.05, 3, "development size n=1e5 f=0.05 nsteps=3"),
}
# batched workloads: name -> (nseries, n, f, nsteps, description)
BATCHED = {
    "c3": (4096, 16384, 0.3, 2, "batched 4096 series x n=16384 f=0.3 nsteps=2 (BASELINE configs[2])"),
    "c3small": (256, 16384, 0.3, 2, "development: 256 series x n=16384 f=0.3 nsteps=2"),
}
FLOP_PER_PAIR = 17.0
# dram__bytes_read.sum + dram__bytes_write.sum of one fit_kernel launch, from the committed
# `ncu --set full` captures (profiles/README.md); None where no capture exists.
NCU_TRAFFIC = {
    "c2": {"bytes": 48.5e6 + 349.2e6, "from": "profiles/r01_fit_kernel_c2_ncu_full.csv (mean of 2 launches; K=10 "
                                              "window split writes 400 MB of partial moments; 0.15% of DRAM peak)"},
}
BYTES_PER_POINT_PASS = 56.0


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS) + sorted(BATCHED))
    ap.add_argument("--method", default="direct", choices=["direct", "tree", "auto"],
                    help="evaluator of a lowest() pass: direct O(n ns) kernel (roofline claim) or moment tree")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: exchange of the robustness step -- kernels over NVLink peer memory, or NCCL")
    ap.add_argument("--no-tree", action="store_true",
                    help="N = 1: skip the extra moment-tree measurements (same workload and n=1e8)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def make_data(n):
    from curvefit_b200 import datagen
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    return x, y


# ----------------------------------------------------------------------------
# CPU arm: the oracle on the host cores, on a bounded sample of output points
# ----------------------------------------------------------------------------
def cpu_sample(x, y, f, nsteps, seconds_target=6.0):
    """Times the oracle's lowest() (all host threads, OpenMP over points) on a
    stratified sample of output points: one unweighted and one weighted pass.
    Returns (points_per_sec for a full smooth, cores, sample description)."""
    from oracle import oracle
    n = x.size
    cores = oracle.max_threads()
    ns = oracle.span(n, f)
    rw = np.ones(n)
    m = max(cores * 32, 64)
    # calibrate m so that one sample pass lasts about seconds_target / 2
    idx = np.linspace(0, n - 1, m).astype(np.int32)
    t0 = time.perf_counter()
    oracle.lowest_at(x, y, f, idx, None)
    dt = time.perf_counter() - t0
    m = int(min(max(m * (seconds_target / 2) / max(dt, 1e-4), m), 200000, n))
    idx = np.linspace(0, n - 1, m).astype(np.int32)
    t0 = time.perf_counter()
    oracle.lowest_at(x, y, f, idx, None)          # pass 1: userw = .false.
    t1 = time.perf_counter()
    oracle.lowest_at(x, y, f, idx, rw)            # later passes: userw = .true.
    t2 = time.perf_counter()
    per_point = ((t1 - t0) + nsteps * (t2 - t1)) / m   # seconds per output point per smooth
    pps = 1.0 / per_point
    ns_per_pair = ((t2 - t0) / 2) / m / ns * cores * 1e9
    desc = (f"{m} stratified output points x 2 passes (1 unweighted, 1 weighted) of n={n} ns={ns}, "
            f"oracle lowest() on {cores} threads, extrapolated to {nsteps + 1} passes; "
            f"{ns_per_pair:.2f} ns/pair/core; |res| sort not included")
    return pps, cores, desc, (t2 - t0)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, f, nsteps, wname = WORKLOADS[args.workload]
    x, y = make_data(n)
    from oracle import oracle
    oracle.build()
    vals, desc, cores = [], "", 1
    target = 3.0
    for _ in range(max(args.warmup, 1)):
        cpu_sample(x, y, f, nsteps, 1.0)
    t_all = 0.0
    for _ in range(args.steps):
        pps, cores, desc, dt = cpu_sample(x, y, f, nsteps, target)
        vals.append(pps)
        t_all += dt
    v = float(np.median(vals))
    line = {
        "impl": "reference", "metric": "LOWESS fitted points/sec", "value": v, "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wname, "n": n, "f": f, "nsteps": nsteps},
        "cpu_baseline": {"value": v, "unit": "points/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "oracle = C restatement of the reference (no Fortran compiler in this image); "
                "ms_per_step is the extrapolated full smooth()",
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------
# clocks sampler
# ----------------------------------------------------------------------------
class Clocks:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self.stop = threading.Event()
        self.th = None

    def _loop(self):
        while not self.stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.samples.append(parts)
            except Exception:
                pass
            self.stop.wait(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._loop, daemon=True)
        self.th.start()
        return self

    def __exit__(self, *a):
        self.stop.set()
        self.th.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        mhz = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for k, nm in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(mhz)}


# ----------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------
def run_ours(args):
    import ctypes as C

    import torch
    import torch.distributed as dist

    from curvefit_b200 import _lib, lowess_smoothing, regression
    from curvefit_b200.distributed import DeviceStages, ShardedLowess, torch_stream_handle

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, f, nsteps, wname = WORKLOADS[args.workload]
    x, y = make_data(n)
    if args.method != "direct":
        os.environ["CF_LOWESS_METHOD"] = args.method     # the stateless e2e calls use it too
    lib = _lib.load()
    ns = lib.cf_lowess_span(n, f)

    stream = torch_stream_handle()
    fit = lowess_smoothing(device=local, stream=stream)
    fit.set_method(args.method)
    fit.initialize(x, y, srt=False)           # resident in HBM from here on
    stages = DeviceStages(fit, rank, world)
    transport = "none"
    if world > 1:
        transport = "p2p" if (args.transport == "p2p" and stages.attach_peers()) else "nccl"
    job = ShardedLowess(stages)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        if world == 1:
            return fit.run_resident(f, nsteps)
        lib.cf_lowess_reset_windows(fit._h)       # a step is a complete smooth(): windows included
        job.run(f, nsteps, gather=True)
        return None

    for _ in range(args.warmup):
        step()
    peak_tf, peak_mhz = regression.fp64_peak(local, 300.0) if rank == 0 else (0.0, 0.0)
    barrier()

    ms_steps, fit_ms, fit_launches, launches = [], 0.0, 0, 0
    with Clocks(local) as clk:
        for _ in range(args.steps):
            flush.fill_(1)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = lib.cf_lowess_launch_count(fit._h)
            e0.record()
            st = step()
            e1.record()
            barrier()
            ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            ms_steps.append(ms)
            if st is not None:
                fit_ms += st.ms_fit
                fit_launches += st.fit_launches
            launches += lib.cf_lowess_launch_count(fit._h) - l0
    total_ms = float(sum(ms_steps))
    value = n * args.steps / (total_ms * 1e-3)
    if world > 1 and stages.p2p and stages.peer_error():
        raise RuntimeError("a peer-memory wait timed out during the timed region")

    line = None
    if rank == 0:
        line = {
            "metric": "LOWESS fitted points/sec", "value": value, "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": wname, "n": n, "f": f, "nsteps": nsteps, "ns": ns,
                       "x": "grid (i-1)*dx", "y": "0.5sin20x+cos5x*exp(-.1x)+U(-.5,.5) splitmix64",
                       "l2": "256 MiB flush write between timed steps",
                       "parallelism": f"range-sharded x{world}" if world > 1 else "single GPU",
                       "method": args.method, "transport": transport},
            "gpu_launches_note": "kernels of libcurvefit_cuda launched by rank 0 inside the timed region",
            "gpu_launches": launches,
            "clocks": clk.summary(),
        }
        if world == 1 and fit_launches and args.method != "direct":
            line["roofline"] = {
                "bound": "fp64", "achieved": None, "peak": peak_tf, "unit": "TFLOP/s", "frac": None,
                "traffic": None, "kernel": "fit stage = per-pass tree rebuild (cf::tree_block_kernel ...) + cf::tree_eval_kernel",
                "avg_launch_ms": fit_ms / fit_launches, "share_of_step": fit_ms / total_ms,
                "note": "moment-tree evaluator: O(n log ns) work, no 17 flop/pair accounting applies; "
                        "ncu: profiles/r01_tree_eval_c4_ncu_full.csv (FP64 pipe 52% active, l1tex 73%). The roofline "
                        "claim of this repo is the direct kernel's (default --method direct)."}
        elif world == 1 and fit_launches:
            avg_ms = fit_ms / fit_launches
            flops = FLOP_PER_PAIR * n * ns
            ach = flops / (avg_ms * 1e-3) / 1e12
            line["roofline"] = {
                "bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": ach / peak_tf if peak_tf else None,
                "traffic": NCU_TRAFFIC.get(args.workload, {}).get("bytes"),
                "traffic_source": NCU_TRAFFIC.get(args.workload, {}).get("from"),
                "kernel": "cf::fit_kernel", "avg_launch_ms": avg_ms,
                "share_of_step": fit_ms / total_ms,
                "peak_source": f"measured here: cf_fp64_peak DFMA microbenchmark ({peak_mhz:.0f} MHz equiv.; "
                               "MEASURED_PEAKS.json has no FP64 figure); nominal 37.2 TFLOP/s at 1965 MHz",
                "algorithmic": "17 flop/pair x n x ns pairs per launch (SURVEY 8d); the kernel issues "
                               "10 FP64 instr/pair so 100% pipe use = 85% of this flop peak",
                "hbm": {"algorithmic_bytes_per_launch": 24.0 * n + 8.0 * n,
                        "gbs": (32.0 * n) / (avg_ms * 1e-3) / 1e9},
            }

    # ---- e2e: host buffers through the stateless C ABI -------------------------------
    if world == 1 and not args.no_e2e:
        xp = torch.from_numpy(x).pin_memory().numpy()
        yp = torch.from_numpy(y).pin_memory().numpy()
        outs = tuple(torch.empty(n, dtype=torch.float64).pin_memory().numpy() for _ in range(3))
        regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)      # warm-up
        reps = max(1, min(args.steps, 3))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / reps
        line["e2e"] = {"value": n / dt, "unit": "points/s", "h2d_bytes_per_step": 16 * n,
                       "d2h_bytes_per_step": 24 * n, "ms_per_step": dt * 1e3,
                       "api": "cf_lowess_smooth(host x, y -> host ys, rw, res), pinned buffers, wall clock"}
    elif world > 1 and not args.no_e2e:
        # every rank uploads the series from pinned host memory, the ranks smooth it together,
        # rank 0 reads ys, rw, res back -- all inside the timed region, max over ranks
        xp = torch.from_numpy(x).pin_memory().numpy()
        yp = torch.from_numpy(y).pin_memory().numpy()
        outs = tuple(torch.empty(n, dtype=torch.float64).pin_memory().numpy() for _ in range(3)) if rank == 0 else None
        job.smooth_host(xp, yp, f, nsteps, outs)
        reps = max(1, min(args.steps, 3))
        barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            job.smooth_host(xp, yp, f, nsteps, outs)
        barrier()
        t = torch.tensor([(time.perf_counter() - t0) / reps], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        if rank == 0:
            line["e2e"] = {"value": n / dt, "unit": "points/s", "h2d_bytes_per_step": 16 * n * world,
                           "d2h_bytes_per_step": 24 * n, "ms_per_step": dt * 1e3,
                           "api": "ShardedLowess.smooth_host(host x, y on every rank -> host ys, rw, res on "
                                  "rank 0), pinned buffers, wall clock, max over ranks"}
    elif rank == 0:
        line["e2e"] = None

    # ---- moment-tree evaluator beside the direct kernel (N = 1 only) -----------------------
    # Own accounting: never part of `value` or of the direct kernel's roofline.
    if world == 1 and args.method == "direct" and not args.no_tree:
        line["tree"] = tree_numbers(args, fit, flush, x, y, f, nsteps)

    # ---- CPU baseline on the host cores (rank 0, N = 1 only) ----------------------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import oracle
        oracle.build()
        pps, cores, desc, _ = cpu_sample(x, y, f, nsteps, 12.0)
        line["cpu_baseline"] = {"value": pps, "unit": "points/s", "cores": cores, "kind": "port",
                                "sample": desc}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        stages.detach_peers()
        dist.barrier()
        dist.destroy_process_group()


def tree_numbers(args, fit, flush, x, y, f, nsteps):
    """The O(n log ns) moment-tree evaluator (csrc/tree.cu) on the bench workload and on the
    BASELINE headline series (n = 1e8, f = 0.05, 3 robustness iterations), device-resident,
    CUDA-event timed per step with the L2 flush in between; plus its largest deviation from
    the direct kernel's ys on the bench workload."""
    import torch

    from curvefit_b200 import lowess_smoothing
    from curvefit_b200.regression import _dp

    def timed(obj, ff, k, reps):
        obj.run_resident(ff, k)
        ms = []
        for _ in range(reps):
            flush.fill_(1)
            torch.cuda.synchronize()
            ms.append(obj.run_resident(ff, k).ms_total)
        return float(np.median(ms))

    out = {}
    ys_direct = np.empty(x.size)
    fit._lib.cf_lowess_fetch(fit._h, _dp(ys_direct), None, None)
    fit.set_method("tree")
    ms = timed(fit, f, nsteps, max(args.steps, 3))
    ys_tree = np.empty(x.size)
    fit._lib.cf_lowess_fetch(fit._h, _dp(ys_tree), None, None)
    fit.set_method("direct")
    out["same_workload"] = {"value": x.size / (ms * 1e-3), "unit": "points/s", "ms_per_step": ms,
                            "max_abs_diff_vs_direct": float(np.max(np.abs(ys_tree - ys_direct))),
                            "max_abs_ys": float(np.max(np.abs(ys_direct)))}
    if not args.no_e2e:
        # the same host-buffer C ABI call as `e2e`, with CF_LOWESS_METHOD=tree
        from curvefit_b200 import regression
        xp = torch.from_numpy(x).pin_memory().numpy()
        yp = torch.from_numpy(y).pin_memory().numpy()
        outs = tuple(torch.empty(x.size, dtype=torch.float64).pin_memory().numpy() for _ in range(3))
        os.environ["CF_LOWESS_METHOD"] = "tree"
        try:
            regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)
            dt = (time.perf_counter() - t0) / 5
        finally:
            os.environ.pop("CF_LOWESS_METHOD", None)
        out["same_workload"]["e2e"] = {"value": x.size / dt, "unit": "points/s", "ms_per_step": dt * 1e3,
                                       "h2d_bytes_per_step": 16 * x.size, "d2h_bytes_per_step": 24 * x.size}
    n4, f4, k4, name4 = WORKLOADS["c4"]
    if x.size != n4:
        x4, y4 = make_data(n4)
        big = lowess_smoothing(device=fit._device, stream=fit._stream)
        big.set_method("tree")
        big.initialize(x4, y4, srt=False)
        ms4 = timed(big, f4, k4, 2)
        out["headline_n1e8"] = {"workload": name4.split(" (")[0], "value": n4 / (ms4 * 1e-3), "unit": "points/s",
                                "ms_per_step": ms4}
        big.close()
    out["note"] = ("moment-tree evaluator, equal to the direct kernel to rounding (tolerance 1e-10); "
                   "separate accounting from `value` / `roofline`")
    return out


# ----------------------------------------------------------------------------
# batched workloads (BASELINE configs[2]): independent series, partitioned over ranks
# ----------------------------------------------------------------------------
def make_batched(nseries, n, first=0):
    from curvefit_b200 import datagen
    x1 = datagen.grid_x(n)
    base = datagen.signal(x1)
    x = np.broadcast_to(x1, (nseries, n)).copy()
    y = np.empty((nseries, n))
    for k in range(nseries):
        y[k] = base + (datagen.uniform01(datagen.SEED_Y + ((first + k) << 32), n) - 0.5)
    return x, y


def run_batched_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    nseries, n, f, nsteps, wname = BATCHED[args.workload]
    from oracle import oracle
    oracle.build()
    cores = oracle.max_threads()
    x, y = make_batched(2, n)
    oracle.lowess_mt(x[0], y[0], f, nsteps)
    vals = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        for k in range(2):
            oracle.lowess_mt(x[k], y[k], f, nsteps)
        vals.append(2 * n / (time.perf_counter() - t0))
    v = float(np.median(vals))
    desc = f"2 of {nseries} series in full (oracle lowess on {cores} threads), scaled"
    print(json.dumps({
        "impl": "reference", "metric": "LOWESS fitted points/sec", "value": v, "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * nseries * n / v, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": wname, "nseries": nseries, "n": n, "f": f, "nsteps": nsteps},
        "cpu_baseline": {"value": v, "unit": "points/s", "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def run_batched(args):
    import torch
    import torch.distributed as dist

    from curvefit_b200 import _lib, regression
    from curvefit_b200.distributed import torch_stream_handle

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nseries, n, f, nsteps, wname = BATCHED[args.workload]
    if args.method != "direct":
        os.environ["CF_LOWESS_METHOD"] = args.method
    mine = nseries // world            # series are independent: partition, no collective
    x, y = make_batched(mine, n, first=rank * mine)
    ns = _lib.load().cf_lowess_span(n, f)
    stream = torch_stream_handle()
    b = regression.batched_lowess(device=local, stream=stream)
    b.set_method(args.method)
    b.initialize(x, y)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        b.run_resident(f, nsteps)
    peak_tf, peak_mhz = regression.fp64_peak(local, 300.0) if rank == 0 else (0.0, 0.0)
    barrier()
    ms_steps, fit_ms, fit_launches, launches = [], 0.0, 0, 0
    with Clocks(local) as clk:
        for _ in range(args.steps):
            flush.fill_(1)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            st = b.run_resident(f, nsteps)
            e1.record()
            barrier()
            ms = e0.elapsed_time(e1)
            if world > 1:
                t = torch.tensor([ms], device="cuda", dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t.item())
            ms_steps.append(ms)
            fit_ms += st.ms_fit
            fit_launches += st.fit_launches
            launches += st.launches
    total_ms = float(sum(ms_steps))
    if rank == 0:
        value = mine * world * n * args.steps / (total_ms * 1e-3)
        avg_ms = fit_ms / max(fit_launches, 1)
        ach = FLOP_PER_PAIR * n * ns * mine / (avg_ms * 1e-3) / 1e12
        line = {
            "metric": "LOWESS fitted points/sec", "value": value, "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": wname, "nseries": nseries, "n": n, "f": f, "nsteps": nsteps, "ns": ns,
                       "l2": "inputs larger than L2 (resident state > 2 GB per GPU) + 256 MiB flush",
                       "parallelism": f"series partitioned x{world}, no collective"},
            "gpu_launches": launches, "clocks": clk.summary(),
            "roofline": {"bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                         "frac": ach / peak_tf if peak_tf else None, "traffic": None,
                         "kernel": "cf::fit_kernel", "avg_launch_ms": avg_ms,
                         "share_of_step": fit_ms / total_ms,
                         "peak_source": f"measured here: cf_fp64_peak ({peak_mhz:.0f} MHz equiv.)"},
        }
        line["config"]["method"] = args.method
        if args.method != "direct":
            line["roofline"].update({"achieved": None, "frac": None,
                                     "kernel": "fit stage = per-pass tree rebuild (cf::tree_block_kernel ...) + cf::tree_eval_kernel",
                                     "note": "moment-tree evaluator: the 17 flop/pair accounting does not apply"})
        if world == 1 and not args.no_e2e:
            k = min(mine, 512)
            xp = torch.from_numpy(x[:k].copy()).pin_memory().numpy()
            yp = torch.from_numpy(y[:k].copy()).pin_memory().numpy()
            regression.cf_lowess_smooth_batched(xp, yp, f, nsteps)
            t0 = time.perf_counter()
            regression.cf_lowess_smooth_batched(xp, yp, f, nsteps)
            dt = time.perf_counter() - t0
            line["e2e"] = {"value": k * n / dt, "unit": "points/s", "h2d_bytes_per_step": 16 * k * n,
                           "d2h_bytes_per_step": 24 * k * n, "ms_per_step": dt * 1e3,
                           "api": f"cf_lowess_smooth_batched on {k} of the series (host buffers, wall clock)"}
        if world == 1 and not args.no_cpu_baseline:
            from oracle import oracle
            oracle.build()
            t0 = time.perf_counter()
            oracle.lowess_mt(x[0], y[0], f, nsteps)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "points/s", "cores": oracle.max_threads(),
                                    "kind": "port", "sample": "1 series in full, oracle lowess on all threads"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    batched = args.workload in BATCHED
    if args.impl == "reference":
        run_batched_reference(args) if batched else run_reference(args)
    else:
        run_batched(args) if batched else run_ours(args)


if __name__ == "__main__":
    main()
