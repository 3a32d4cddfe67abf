#!/usr/bin/env python
"""bench.py -- LOWESS fitted points/sec on B200 (BASELINE.json metric).

A step = one complete smooth() (windows + nsteps+1 fit passes + nsteps robustness
updates) of the workload, inputs resident in HBM.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    torchrun ... bench.py --gpus N ...        (one rank per GPU, N > 1)

The default invocation prints ONE compact JSON line (< 3 kB; what the strings used to explain
is in DESIGN.md section 6) that carries, at every N:

  value / ms_per_step / stages / e2e / roofline / vs_single_gpu
                           the configuration BASELINE.json's metric is quoted on: n = 1e8,
                           f = 0.05, nsteps = 3, range-sharded over the N ranks, on the
                           moment-tree evaluator (the product's fastest parity-checked path);
                           roofline = HBM, 56 B/point/pass against MEASURED_PEAKS.json
  c2_direct                BASELINE configs[1] (n = 1e6, f = 0.05, nsteps = 3) on the DIRECT
                           kernel, a complete run: value, e2e, and the FP64 roofline of
                           cf::fit_kernel -- the O(n ns) sweep north_star's roofline claim is on
  c4_direct_sample         ONE direct-kernel pass at the headline geometry (ns = 5e6; 5.5 min in
                           full) over two waves of strided 512-point tiles per rank, with its own
                           roofline fraction and an oracle check at sampled points
  cpu_baseline (N = 1)     the oracle (C restatement of the reference) on the host cores, on a
                           bounded sample of output points of the n = 1e8 series

`--workload c2|c4|c5|c3|...` runs a single workload instead (with `--method direct|tree`).
`--impl reference`: the reference's CPU algorithm (the oracle; the Fortran reference cannot be
compiled in this image) on all host cores AND on one thread, on a bounded sample.
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
    "c4": (100_000_000, 0.05, 3, "single series n=1e8 f=0.05 nsteps=3 (BASELINE headline)"),
    "small": (100_000, 0.05, 3, "development size n=1e5 f=0.05 nsteps=3"),
    "mid": (4_000_000, 0.05, 3, "development size n=4e6 f=0.05 nsteps=3"),
}
# batched workloads: name -> (nseries, n, f, nsteps, description)
BATCHED = {
    "c3": (4096, 16384, 0.3, 2, "batched 4096 series x n=16384 f=0.3 nsteps=2 (BASELINE configs[2])"),
    "c3small": (256, 16384, 0.3, 2, "development: 256 series x n=16384 f=0.3 nsteps=2"),
}
FLOP_PER_PAIR = 17.0
TILE = 512                # output points per tile of the direct kernel
# dram__bytes_read.sum + dram__bytes_write.sum of one fit_kernel launch, from the committed
# `ncu --set full` captures (profiles/README.md); None where no capture exists.
NCU_TRAFFIC = {
    "c2": {"bytes": 42.7e6 + 347.9e6, "from": "profiles/r02_fit_kernel_c2_ncu_full.csv (one launch of this code; the K=10 "
                                              "window split writes 348 MB of partial moments; 0.15% of DRAM peak)"},
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="default", choices=["default"] + sorted(WORKLOADS) + sorted(BATCHED))
    ap.add_argument("--method", default=None, choices=["direct", "tree", "auto"],
                    help="evaluator of a lowest() pass (single-workload runs): direct O(n ns) kernel or moment tree")
    ap.add_argument("--transport", default="p2p", choices=["p2p", "nccl"],
                    help="N > 1: exchange of the robustness step -- kernels over NVLink peer memory, or NCCL")
    ap.add_argument("--no-headline", action="store_true", help="default run: skip the n = 1e8 part")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def make_data(n):
    from curvefit_b200 import datagen
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    return x, y


def series_config(name, method, world, transport):
    """The `config` object of a single-series workload; both arms print the same one."""
    n, f, nsteps, wname = WORKLOADS[name]
    ns = max(min(int(f * float(n)), n), 2)
    return {"workload": wname, "n": n, "f": f, "nsteps": nsteps, "ns": ns,
            "x": "grid (i-1)*dx", "y": "0.5sin20x+cos5x*exp(-.1x)+U(-.5,.5) splitmix64",
            "l2": "256 MiB flush write between timed steps",
            "parallelism": f"range-sharded x{world}" if world > 1 else "single GPU",
            "method": method, "transport": transport if world > 1 else "none"}


# ----------------------------------------------------------------------------
# CPU arm: the oracle on the host cores, on a bounded sample of output points
# ----------------------------------------------------------------------------
def cpu_sample(x, y, f, nsteps, threads, seconds_target=6.0, robust=True):
    """Times the oracle's lowest() on `threads` host threads (OpenMP over points; set
    explicitly -- launchers export OMP_NUM_THREADS=1) over a stratified sample of output
    points, one unweighted and one weighted pass, and the robustness step (|res| full sort +
    bisquare, single-threaded like the reference) in full, and extrapolates one smooth():
        t = n * (t_unweighted + nsteps * t_weighted) / m + nsteps * t_robust."""
    from oracle import oracle
    n = x.size
    threads = oracle.set_threads(threads)
    ns = oracle.span(n, f)
    rw = np.ones(n)
    m = max(threads * 8, 16)
    idx = np.linspace(0, n - 1, m).astype(np.int32)
    t0 = time.perf_counter()
    oracle.lowest_at(x, y, f, idx, None)           # calibration (also warms the caches)
    dt = time.perf_counter() - t0
    m = int(min(max(m * (seconds_target / 2) / max(dt, 1e-4), m), 200000, n))
    idx = np.linspace(0, n - 1, m).astype(np.int32)
    t0 = time.perf_counter()
    ya = oracle.lowest_at(x, y, f, idx, None)      # pass 1: userw = .false.
    t1 = time.perf_counter()
    oracle.lowest_at(x, y, f, idx, rw)             # later passes: userw = .true.
    t2 = time.perf_counter()
    t_rob = t_rob_measured = 0.0
    if robust and nsteps > 0:
        res = y - np.interp(x, x[idx], ya["ys"])   # residual-like values; only their sort cost matters
        nr = min(n, 4_000_000)                     # bounded: n log n scaling applied above that
        t3 = time.perf_counter()
        oracle.robust_weights(res[:nr])
        t_rob = t_rob_measured = time.perf_counter() - t3
        if nr < n:
            t_rob *= (n * np.log2(n)) / (nr * np.log2(nr))
    fit_s = n * ((t1 - t0) + nsteps * (t2 - t1)) / m
    total_s = fit_s + nsteps * t_rob
    ns_per_pair = ((t2 - t0) / 2) / m / ns * threads * 1e9
    desc = (f"{m} stratified output points x 2 passes (1 unweighted, 1 weighted) of n={n} ns={ns}, oracle "
            f"lowest() on {threads} thread(s), extrapolated to {nsteps + 1} passes; {ns_per_pair:.2f} ns/pair/core; "
            f"+ {nsteps} x the |res| sort + bisquare step ({t_rob * 1e3:.0f} ms each, single thread"
            f"{', scaled n log n from 4e6' if robust and n > 4_000_000 else ''})")
    return {"value": n / total_s, "cores": threads, "sample": desc, "measured_s": (t2 - t0) + t_rob_measured,
            "extrapolated_smooth_s": total_s, "ns_per_pair_per_core": ns_per_pair}


def run_reference_default(args):
    """CPU arm of the default line: the reference's algorithm (the oracle) on the n = 1e8 series.
    A full smooth is 5e14 (point, element) pairs -- weeks on the host cores -- so every step is a
    bounded sample of output points, extrapolated (cpu_sample); beside it, BASELINE configs[1]
    (n = 1e6) is run IN FULL once, all four passes, as a measured like-for-like for `c2_direct`."""
    from oracle import oracle
    oracle.build()
    cores = oracle.host_cores()
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    n, f, nsteps, _ = WORKLOADS["c4"]
    x, y = make_data(n)
    for _ in range(max(min(args.warmup, 2), 1)):
        cpu_sample(x, y, f, nsteps, cores, 0.5, robust=False)
    t_wall = time.perf_counter()
    runs = [cpu_sample(x, y, f, nsteps, cores, 3.0, robust=False) for _ in range(max(1, min(args.steps, 5)))]
    runs.sort(key=lambda r: r["value"])
    allc = runs[len(runs) // 2]
    one = cpu_sample(x, y, f, nsteps, 1, 4.0, robust=False)
    wall = time.perf_counter() - t_wall
    v = allc["value"]
    line = {
        "impl": "reference", "metric": "LOWESS fitted points/sec", "value": v, "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": series_config("c4", "tree", world, args.transport),
        "cpu_baseline": {"value": v, "unit": "points/s", "cores": allc["cores"], "kind": "port",
                         "sample": allc["sample"][:200]},
        "single_thread": {"value": one["value"], "cores": 1},
        "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "timed_region": {"measured_sample_s_per_step": allc["measured_s"], "wall_s_all_steps": wall,
                         "extrapolated": True},
        "gpu_launches": 0,
        "fortran_reference": "not built: no Fortran compiler (oracle/ref/build_ref.sh)",
    }
    if not args.no_headline:
        n2, f2, k2, _ = WORKLOADS["c2"]
        x2, y2 = make_data(n2)
        oracle.set_threads(cores)
        t0 = time.perf_counter()
        oracle.lowess_mt(x2, y2, f2, k2)
        dt = time.perf_counter() - t0
        line["c2_full"] = {"n": n2, "f": f2, "nsteps": k2, "value": n2 / dt, "seconds": dt, "cores": cores,
                           "extrapolated": False}
    print(json.dumps(round_floats(line), separators=(",", ":")), flush=True)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    if args.workload == "default":
        return run_reference_default(args)
    name = args.workload
    n, f, nsteps, wname = WORKLOADS[name]
    x, y = make_data(n)
    from oracle import oracle
    oracle.build()
    cores = oracle.host_cores()
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    method = args.method or "direct"
    for _ in range(max(min(args.warmup, 2), 1)):
        cpu_sample(x, y, f, nsteps, cores, 0.5, robust=False)
    t_wall = time.perf_counter()
    runs = [cpu_sample(x, y, f, nsteps, cores, 3.0) for _ in range(max(1, min(args.steps, 5)))]
    runs.sort(key=lambda r: r["value"])
    allc = runs[len(runs) // 2]
    one = cpu_sample(x, y, f, nsteps, 1, 4.0)
    wall = time.perf_counter() - t_wall
    v = allc["value"]
    line = {
        "impl": "reference", "metric": "LOWESS fitted points/sec", "value": v, "unit": "points/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * n / v, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": series_config(name, method, world, args.transport),
        "cpu_baseline": {"value": v, "unit": "points/s", "cores": allc["cores"], "kind": "port",
                         "sample": allc["sample"]},
        "single_thread": {"value": one["value"], "unit": "points/s", "cores": 1, "sample": one["sample"],
                          "note": "the reference's real execution model: no OpenMP anywhere (CMakeLists.txt:114)"},
        "e2e": {"value": v, "unit": "points/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "timed_region": {"measured_sample_s_per_step": allc["measured_s"], "wall_s_all_steps": wall,
                         "ms_per_step_is": "EXTRAPOLATED full smooth() = n/m x sample passes + nsteps x |res| sort"},
        "gpu_launches": 0,
        "note": "oracle = C restatement of the reference (no Fortran compiler in this image); value = all host "
                "cores (OpenMP over output points, thread count set explicitly), single_thread beside it",
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
class Env:
    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from curvefit_b200 import _lib
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.host_group = None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            # host-only rendezvous: an NCCL barrier is a kernel that SPINS on every waiting rank's GPU,
            # which must stay idle while rank 0 alone drives all devices (one-process C-ABI arm)
            self.host_group = dist.new_group(backend="gloo")
        self.lib = _lib.load()
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2
        self.args = args

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def host_barrier(self):
        if self.world > 1:
            self.torch.cuda.synchronize()
            self.dist.barrier(group=self.host_group)

    def max_over_ranks(self, vals):
        if self.world == 1:
            return [float(v) for v in vals]
        t = self.torch.tensor(list(vals), device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def pinned(self, a):
        return self.torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()

    def pinned_empty(self, n):
        return self.torch.empty(n, dtype=self.torch.float64).pin_memory().numpy()


def measure_series(env, name, method, steps, warmup, want_e2e, clocks=None, sample_direct=False,
                   compare_single=False):
    """One single-series workload, range-sharded over the ranks: device-resident steps (CUDA
    events, L2 flush between steps, max over ranks), per-stage times, optional e2e with host
    buffers, optional comparison with a single-GPU run, optional sampled direct-kernel pass."""
    import ctypes as C

    from curvefit_b200 import _lib, lowess_smoothing, regression
    from curvefit_b200.distributed import DeviceStages, ShardedLowess, torch_stream_handle
    torch, lib, args = env.torch, env.lib, env.args
    world, rank = env.world, env.rank
    n, f, nsteps, wname = WORKLOADS[name]
    x, y = make_data(n)
    ns = lib.cf_lowess_span(n, f)
    stream = torch_stream_handle()
    fit = lowess_smoothing(device=env.local, stream=stream)
    fit.set_method(method)
    fit.initialize(x, y, srt=False)           # resident in HBM from here on
    stages = DeviceStages(fit, rank, world)
    transport = "none"
    if world > 1:
        transport = "p2p" if (args.transport == "p2p" and stages.attach_peers()) else "nccl"
    job = ShardedLowess(stages)

    def step():
        if world == 1:
            return fit.run_resident(f, nsteps)
        lib.cf_lowess_reset_windows(fit._h)       # a step is a complete smooth(): windows included
        job.run(f, nsteps, gather=False)          # every rank ends up holding its block of ys, rw, res
        return job.last_stats

    for _ in range(warmup):
        step()
    env.barrier()
    ms_steps, launches = [], 0
    stage = {"windows": 0.0, "fit": 0.0, "robust": 0.0, "tree_build": 0.0}
    fit_launches = 0
    for _ in range(steps):
        env.flush.fill_(1)
        env.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = lib.cf_lowess_launch_count(fit._h)
        e0.record()
        st = step()
        e1.record()
        env.barrier()
        vals = [e0.elapsed_time(e1)]
        if st is not None:
            vals += [st.ms_windows, st.ms_fit, st.ms_robust, st.ms_tree_build]
        vals = env.max_over_ranks(vals)
        ms_steps.append(vals[0])
        if st is not None:
            for k, v in zip(("windows", "fit", "robust", "tree_build"), vals[1:]):
                stage[k] += v
            fit_launches += st.fit_launches
        launches += lib.cf_lowess_launch_count(fit._h) - l0
    total_ms = float(sum(ms_steps))
    if world > 1:
        job.check_all()
    out = {"workload": wname, "method": method, "transport": transport,
           "value": n * steps / (total_ms * 1e-3), "unit": "points/s", "n_gpus": world,
           "ms_per_step": total_ms / steps, "steps": steps, "warmup": warmup,
           "ms_min_max": [min(ms_steps), max(ms_steps)],
           "stages_ms_per_step": {k: v / steps for k, v in stage.items()} if fit_launches else None,
           "stages_note": "CUDA events inside cf_lowess_run, max over ranks; tree_build is part of fit; "
                          "result left sharded (each rank holds its block of ys, rw, res)",
           "gpu_launches": launches, "fit_launches": fit_launches, "ns": ns}
    out["_fit_ms"] = stage["fit"]

    ys_solo = None
    # ---- the sharded result against a single-GPU run of the same series (outside the timed region)
    if world > 1 and compare_single:
        job.gather_results(with_rw=True)
        torch.cuda.synchronize()
        if rank == 0:
            solo = lowess_smoothing(device=env.local)
            solo.set_method(method)
            solo.initialize(x, y, srt=False)
            t0 = time.perf_counter()
            solo.run_resident(f, nsteps)
            st1 = solo.run_resident(f, nsteps)
            ys1, ysN = np.empty(n), np.empty(n)
            lib.cf_lowess_fetch(solo._h, regression._dp(ys1), None, None)
            lib.cf_lowess_fetch(fit._h, regression._dp(ysN), None, None)
            d = float(np.max(np.abs(ys1 - ysN)))
            out["vs_single_gpu"] = {"max_abs_diff_ys": d, "bitwise_equal_ys": bool(np.array_equal(ys1, ysN)),
                                    "max_abs_ys": float(np.max(np.abs(ys1))),
                                    "single_gpu_ms_same_box": float(st1.ms_total),
                                    "speedup_vs_single_gpu_same_box": float(st1.ms_total) / out["ms_per_step"]}
            assert d <= 1e-12 * max(out["vs_single_gpu"]["max_abs_ys"], 1.0), f"sharded ys differs by {d}"
            solo.close()
            ys_solo = ys1
            del ysN
        env.barrier()

    # ---- e2e: HOST buffers in, HOST buffers out, copies inside the timed region ---------------
    if want_e2e and world == 1:
        xp, yp = env.pinned(x), env.pinned(y)
        outs = tuple(env.pinned_empty(n) for _ in range(3))
        regression.set_default_method(method)
        try:
            regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)      # warm-up
            reps = max(1, min(steps, 3))
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                regression.cf_lowess_smooth(xp, yp, f, nsteps, out=outs)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / reps
        finally:
            regression.set_default_method("direct")
        out["e2e"] = {"value": n / dt, "unit": "points/s", "h2d_bytes_per_step": 16 * n,
                      "d2h_bytes_per_step": 24 * n, "ms_per_step": dt * 1e3,
                      "api": "cf_lowess_smooth(host x, y -> host ys, rw, res), pinned buffers, wall clock"}
        # the overlapped host-buffer path (chunked upload, last pass in pieces) against the resident
        # run of the timed steps above: same series, same evaluator -> the same bits
        same = True
        back = np.empty(n)
        for k, o in enumerate(outs):
            lib.cf_lowess_fetch(fit._h, *[regression._dp(back) if j == k else None for j in range(3)])
            same = same and bool(np.array_equal(back, o))
        out["e2e"]["bitwise_equal_resident"] = same
        assert same, "cf_lowess_smooth (host buffers) differs from the device-resident run"
        del xp, yp, outs, back
    elif want_e2e:
        from curvefit_b200.distributed import SharedHostArray
        xp, yp = env.pinned(x), env.pinned(y)
        # one host result all ranks map (POSIX shared memory); where /dev/shm is too small for it the
        # older form is timed instead: every rank uploads the series, rank 0 reads everything back
        room = [False]
        if rank == 0:
            try:
                fs = os.statvfs("/dev/shm")
                room[0] = fs.f_bavail * fs.f_frsize > 3 * 8 * n + (64 << 20)
            except OSError:
                room[0] = False
        env.dist.broadcast_object_list(room, src=0, group=env.host_group)
        shared_out = bool(room[0])
        shared = [SharedHostArray(n) for _ in range(3)] if shared_out else []      # ys, rw, res
        if shared_out:
            outs = tuple(a.array for a in shared)
        else:
            outs = tuple(env.pinned_empty(n) for _ in range(3)) if rank == 0 else None
        job.smooth_host(xp, yp, f, nsteps, outs, shared_out=shared_out)
        reps = max(1, min(steps, 3))
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            job.smooth_host(xp, yp, f, nsteps, outs, shared_out=shared_out)
        env.barrier()
        dt = env.max_over_ranks([(time.perf_counter() - t0) / reps])[0]
        up = torch.tensor([float(job.last_upload)], device="cuda", dtype=torch.float64)
        env.dist.all_reduce(up)
        out["e2e"] = {"value": n / dt, "unit": "points/s", "h2d_bytes_per_step": int(16 * up.item()),
                      "d2h_bytes_per_step": 24 * n, "ms_per_step": dt * 1e3,
                      "api": ("ShardedLowess.smooth_host(host x, y -> host ys, rw, res in shared page-locked memory): "
                              "each rank uploads its block + halo and writes its own block of the "
                              "result; wall clock incl. the closing barrier, max over ranks") if shared_out else
                             ("ShardedLowess.smooth_host(host x, y on every rank -> host ys, rw, res on rank 0; "
                              "/dev/shm too small for the shared result), pinned buffers, wall clock, max over ranks")}
        if rank == 0 and ys_solo is not None:                   # the assembled host result vs one GPU
            d = float(np.max(np.abs(outs[0] - ys_solo)))
            out["e2e"]["max_abs_diff_ys_vs_single_gpu"] = d
            assert d <= 1e-12 * max(float(np.max(np.abs(ys_solo))), 1.0), f"e2e ys differs by {d}"
        env.barrier()
        for a in shared:
            a.close()
        del xp, yp, outs, shared
        # the same series through the C ABI from ONE process: cf_lowess_smooth_multi on rank 0
        # drives all `world` devices (one host thread each) while the other ranks wait ON THE HOST
        env.host_barrier()
        if rank == 0:
            xp, yp = env.pinned(x), env.pinned(y)
            outs = tuple(env.pinned_empty(n) for _ in range(3))
            devs = (C.c_int32 * world)(*range(world))
            regression.set_default_method(method)
            try:
                call = lambda: lib.cf_lowess_smooth_multi(regression._dp(xp), regression._dp(yp), n, float(f),
                                                          int(nsteps), devs, world, *(regression._dp(o) for o in outs))
                rc = call()                                                   # builds and attaches the handles
                if rc == 0:
                    reps = max(1, min(steps, 3))
                    t0 = time.perf_counter()
                    for _ in range(reps):
                        rc = rc or call()
                    dt1 = (time.perf_counter() - t0) / reps
                if rc == 0 and ys_solo is not None:
                    d = float(np.max(np.abs(outs[0] - ys_solo)))
                    assert d <= 1e-12 * max(float(np.max(np.abs(ys_solo))), 1.0), f"cf_lowess_smooth_multi ys differs by {d}"
                if rc == 0:
                    out["e2e_c_abi_one_process"] = {
                        "value": n / dt1, "unit": "points/s", "ms_per_step": dt1 * 1e3, "devices": world,
                        "api": "cf_lowess_smooth_multi(host x, y -> host ys, rw, res) from rank 0's process alone, "
                               "one host thread per device, pinned buffers, wall clock"}
                else:
                    out["e2e_c_abi_one_process"] = {"error": _lib.last_error()}
            finally:
                regression.set_default_method("direct")
                lib.cf_lowess_multi_release()
            del xp, yp, outs
        env.host_barrier()
    else:
        out["e2e"] = None

    # ---- one DIRECT-kernel pass at this geometry on a strided sample of tiles ---------------------
    if sample_direct:
        out["_direct_sample"] = direct_sample(env, fit, stages, x, y, n, f, ns)

    if world > 1:
        stages.detach_peers()
    fit.close()
    return out


def direct_sample(env, fit, stages, x, y, n, f, ns):
    """Every rank runs ONE weighted pass of the direct kernel (K = 1, as a full pass at this size
    would) over two waves of 512-point tiles spread evenly over its shard, with the robustness
    weights the preceding smooth left on the device.  Rate and roofline fraction come from the
    slowest rank; rank 0 checks sampled points against the oracle's lowest()."""
    from curvefit_b200 import regression
    world, rank = env.world, env.rank
    begin, end = stages.begin, stages.end
    tiles_shard = max((end - begin + TILE - 1) // TILE, 1)
    ntiles = min(2 * 148 * 4, tiles_shard)
    stride = max(tiles_shard // ntiles, 1)
    peak_tf, peak_mhz = regression.fp64_peak(env.local, 300.0)
    env.barrier()
    ms, idx, ys = fit.fit_sample(f, True, stride // 2, stride, ntiles)
    vals = env.max_over_ranks([ms, -peak_tf])
    ms_max, peak_min = vals[0], -vals[1]
    pts = ntiles * TILE                           # per rank
    tf_gpu = FLOP_PER_PAIR * pts * ns / (ms_max * 1e-3) / 1e12
    full_pass_s = (tiles_shard / ntiles) * ms_max * 1e-3
    res = {"kernel": "cf::fit_kernel (K = 1, no window split)", "n": n, "ns": ns, "n_gpus": world,
           "tiles_per_rank": ntiles, "tile_stride": stride, "points_per_rank": pts, "ms": ms_max,
           "roofline": {"bound": "fp64", "achieved": tf_gpu, "peak": peak_min, "unit": "TFLOP/s",
                        "frac": tf_gpu / peak_min if peak_min else None, "traffic": None,
                        "algorithmic": "17 flop/pair x sampled points x ns, per GPU, slowest rank",
                        "peak_source": "cf_fp64_peak DFMA microbenchmark, min over ranks"},
           "points_per_s_all_gpus": pts * world / (ms_max * 1e-3),
           "extrapolated": {"one_pass_s": full_pass_s, "smooth_s_nsteps3": 4 * full_pass_s,
                            "points_per_s_smooth": n / (4 * full_pass_s),
                            "note": "every tile costs the same (ns + 512 staged elements x 512 points), so a full pass "
                                    "= tiles/sampled tiles x this launch; the O(n) stages add < 0.01 %"}}
    if rank == 0:
        from oracle import oracle
        oracle.build()
        oracle.set_threads(oracle.host_cores())
        rw = np.empty(n)
        env.lib.cf_lowess_fetch(fit._h, None, regression._dp(rw), None)
        pick = np.unique(np.linspace(0, idx.size - 1, 24).astype(np.int64))
        t = oracle.lowest_at(x, y, f, idx[pick].astype(np.int32), rw)
        d = np.abs(ys[pick] - t["ys"])
        scale = float(np.max(np.abs(t["ys"])))
        res["oracle_check"] = {"points": int(pick.size), "max_abs_diff": float(d.max()),
                               "max_rel_to_scale": float(d.max() / scale), "tolerance": 1e-10,
                               "ok": bool(d.max() <= 1e-10 * scale)}
        assert res["oracle_check"]["ok"], res["oracle_check"]
    return res


HBM_BYTES_PER_POINT_PASS = 56.0   # x, y, rw read, ys written + the robustness step's share
# one cf::tree_eval_kernel launch at n = 1e8, ns = 5e6 (profiles/r02_tree_eval_c4_ncu_full.csv)
NCU_TREE_EVAL = {"dram_bytes": 15.9e9 + 0.8e9, "fp64_pipe_pct": 53.9}


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback"


def tree_roofline(prim, n, nsteps, world):
    """The tree path against the only bound an O(n log ns) method has: the bytes a pass must move."""
    passes = nsteps + 1
    fit_ms = prim["_fit_ms"] / prim["steps"]                       # all passes of one smooth, slowest rank
    build_ms = prim["stages_ms_per_step"]["tree_build"]
    eval_ms = (fit_ms - build_ms) / passes                         # one cf::tree_eval_kernel launch
    peak, src = hbm_peak()
    ach = HBM_BYTES_PER_POINT_PASS * (n / world) / (eval_ms * 1e-3) / 1e9
    smooth = HBM_BYTES_PER_POINT_PASS * (n / world) * passes / (prim["ms_per_step"] * 1e-3) / 1e9
    return {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": NCU_TREE_EVAL["dram_bytes"] if world == 1 else None,
            "kernel": "cf::tree_eval_kernel", "avg_launch_ms": eval_ms, "per_gpu": True,
            "share_of_step": eval_ms * passes / prim["ms_per_step"],
            "whole_smooth_gbs": smooth, "whole_smooth_frac": smooth / peak,
            "fp64_pipe_pct_ncu": NCU_TREE_EVAL["fp64_pipe_pct"] if world == 1 else None,
            "peak_source": src}


def compact(d, drop=("api", "stages_note", "note", "sample", "algorithmic", "peak_source", "traffic_source",
                     "workload", "unit", "steps", "warmup", "transport", "x", "y", "l2")):
    """Nested objects keep numbers only (the driver keeps a 3 kB tail of the line)."""
    if isinstance(d, dict):
        return {k: compact(v, drop) for k, v in d.items() if k not in drop and not k.startswith("_")}
    if isinstance(d, float):
        return float(f"{d:.6g}")
    return d


def round_floats(d, digits=7):
    if isinstance(d, dict):
        return {k: round_floats(v, digits) for k, v in d.items()}
    if isinstance(d, list):
        return [round_floats(v, digits) for v in d]
    if isinstance(d, float):
        return float(f"{d:.{digits}g}")
    return d


def run_default(args):
    """The driver's line: top level = the metric's own configuration (n = 1e8, tree), nested =
    the direct kernel's evidence (C2 in full, one sampled pass at ns = 5e6)."""
    from curvefit_b200 import regression
    env = Env(args)
    rank, world = env.rank, env.world
    n, f, nsteps, wname = WORKLOADS["c4"]
    peak_tf, peak_mhz = regression.fp64_peak(env.local, 300.0)
    peak_tf = -env.max_over_ranks([-peak_tf])[0]
    with Clocks(env.local) as clk:
        head = measure_series(env, "c4", "tree", args.steps, args.warmup, not args.no_e2e,
                              sample_direct=not args.no_headline, compare_single=True)
    line = None
    if rank == 0:
        line = {
            "metric": "LOWESS fitted points/sec", "value": head["value"], "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": head["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": series_config("c4", "tree", world, args.transport),
            "transport_used": head["transport"], "gpu_launches": head["gpu_launches"],
            "clocks": clk.summary(), "ms_min_max": head["ms_min_max"],
            "stages_ms_per_step": head["stages_ms_per_step"],
            "e2e": compact(head["e2e"], drop=("api",)),
            "roofline": tree_roofline(head, n, nsteps, world),
        }
        for k in ("vs_single_gpu", "e2e_c_abi_one_process"):
            if k in head:
                line[k] = compact(head[k])
        if "_direct_sample" in head:
            ds = head["_direct_sample"]
            line["c4_direct_sample"] = compact({
                "tiles_per_rank": ds["tiles_per_rank"], "points_per_rank": ds["points_per_rank"], "ms": ds["ms"],
                "tflops": ds["roofline"]["achieved"], "peak": ds["roofline"]["peak"], "frac": ds["roofline"]["frac"],
                "one_pass_s": ds["extrapolated"]["one_pass_s"], "oracle_check": ds.get("oracle_check")})
    # ---- the direct kernel in full: BASELINE configs[1] ------------------------------------------
    if not args.no_headline:
        n2, f2, k2, _ = WORKLOADS["c2"]
        c2 = measure_series(env, "c2", "direct", args.steps, args.warmup, not args.no_e2e, compare_single=True)
        if rank == 0:
            avg_ms = c2["_fit_ms"] / max(c2["fit_launches"], 1)
            ach = FLOP_PER_PAIR * (n2 / world) * c2["ns"] / (avg_ms * 1e-3) / 1e12
            line["c2_direct"] = compact({
                "n": n2, "f": f2, "nsteps": k2, "ns": c2["ns"], "value": c2["value"], "ms_per_step": c2["ms_per_step"],
                "stages_ms_per_step": c2["stages_ms_per_step"], "gpu_launches": c2["gpu_launches"],
                "e2e": c2["e2e"], "vs_single_gpu": c2.get("vs_single_gpu"),
                "roofline": {"bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                             "frac": ach / peak_tf if peak_tf else None,
                             "traffic": NCU_TRAFFIC["c2"]["bytes"] if world == 1 else None,
                             "kernel": "cf::fit_kernel", "avg_launch_ms": avg_ms,
                             "share_of_step": c2["_fit_ms"] / (c2["ms_per_step"] * c2["steps"])}})
    # ---- CPU baseline on the host cores (rank 0, N = 1 only): bounded sample of the n = 1e8 series
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import oracle
        oracle.build()
        x, y = make_data(n)
        allc = cpu_sample(x, y, f, nsteps, oracle.host_cores(), 12.0, robust=False)
        one = cpu_sample(x, y, f, nsteps, 1, 5.0, robust=False)
        line["cpu_baseline"] = {"value": allc["value"], "unit": "points/s", "cores": allc["cores"], "kind": "port",
                                "sample": allc["sample"][:160],
                                "single_thread": {"value": one["value"], "cores": 1}}
        if "c2_direct" in line:
            x2, y2 = make_data(WORKLOADS["c2"][0])
            c = cpu_sample(x2, y2, WORKLOADS["c2"][1], WORKLOADS["c2"][2], oracle.host_cores(), 6.0)
            line["c2_direct"]["cpu_baseline"] = compact({"value": c["value"], "cores": c["cores"], "kind": "port"})
    if rank == 0:
        print(json.dumps(round_floats(line), separators=(",", ":")), flush=True)
    if world > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()


def run_ours(args):
    if args.workload == "default":
        return run_default(args)
    from curvefit_b200 import regression
    env = Env(args)
    rank, world = env.rank, env.world
    single = True
    name = args.workload
    method = args.method or ("direct" if name != "c4" else "tree")
    n, f, nsteps, wname = WORKLOADS[name]
    peak_tf, peak_mhz = regression.fp64_peak(env.local, 300.0)
    peak_tf = -env.max_over_ranks([-peak_tf])[0]

    with Clocks(env.local) as clk:
        prim = measure_series(env, name, method, args.steps, args.warmup, not args.no_e2e,
                              sample_direct=single and name == "c4" and method != "direct",
                              compare_single=True)
    line = None
    if rank == 0:
        ns = prim["ns"]
        line = {
            "metric": "LOWESS fitted points/sec", "value": prim["value"], "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": prim["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": series_config(name, method, world, args.transport),
            "transport_used": prim["transport"],
            "gpu_launches_note": "kernels of libcurvefit_cuda launched by rank 0 inside the timed region",
            "gpu_launches": prim["gpu_launches"],
            "clocks": clk.summary(),
            "stages_ms_per_step": prim["stages_ms_per_step"],
            "e2e": prim["e2e"],
        }
        if "vs_single_gpu" in prim:
            line["vs_single_gpu"] = prim["vs_single_gpu"]
        if prim["fit_launches"] and method == "direct":
            avg_ms = prim["_fit_ms"] / prim["fit_launches"]
            flops = FLOP_PER_PAIR * (n / world) * ns          # per GPU: its shard of output points
            ach = flops / (avg_ms * 1e-3) / 1e12
            line["roofline"] = {
                "bound": "fp64", "achieved": ach, "peak": peak_tf, "unit": "TFLOP/s",
                "frac": ach / peak_tf if peak_tf else None,
                "traffic": NCU_TRAFFIC.get(name, {}).get("bytes") if world == 1 else None,
                "traffic_source": NCU_TRAFFIC.get(name, {}).get("from") if world == 1 else None,
                "kernel": "cf::fit_kernel", "avg_launch_ms": avg_ms, "per_gpu": True,
                "share_of_step": prim["_fit_ms"] / (prim["ms_per_step"] * prim["steps"]),
                "peak_source": f"measured here: cf_fp64_peak DFMA microbenchmark ({peak_mhz:.0f} MHz equiv.; "
                               "MEASURED_PEAKS.json has no FP64 figure), min over ranks; nominal 37.2 TFLOP/s at 1965 MHz",
                "algorithmic": "17 flop/pair x (n / n_gpus) x ns pairs per launch (SURVEY 8d); the kernel issues "
                               "10 FP64 instr/pair so 100% pipe use = 85% of this flop peak; avg_launch_ms is the "
                               "slowest rank's (finalize of the window split included)",
                "hbm": {"algorithmic_bytes_per_launch": 32.0 * n / world,
                        "gbs": (32.0 * n / world) / (avg_ms * 1e-3) / 1e9},
            }
        elif prim["fit_launches"]:
            line["roofline"] = {
                "bound": "fp64", "achieved": None, "peak": peak_tf, "unit": "TFLOP/s", "frac": None, "traffic": None,
                "kernel": "fit stage = per-pass tree rebuild (cf::tree_block_kernel ...) + cf::tree_eval_kernel",
                "avg_launch_ms": prim["_fit_ms"] / prim["fit_launches"],
                "note": "moment-tree evaluator: O(n log ns) work, the 17 flop/pair accounting does not apply; the "
                        "roofline claim of this repo is the direct kernel's (c4_direct_sample, or --method direct)"}
        if "_direct_sample" in prim:
            line["c4_direct_sample"] = prim["_direct_sample"]

    # ---- CPU baseline on the host cores (rank 0, N = 1 only) ----------------------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import oracle
        oracle.build()
        x, y = make_data(n)
        allc = cpu_sample(x, y, f, nsteps, oracle.host_cores(), 10.0)
        one = cpu_sample(x, y, f, nsteps, 1, 5.0)
        line["cpu_baseline"] = {"value": allc["value"], "unit": "points/s", "cores": allc["cores"], "kind": "port",
                                "sample": allc["sample"],
                                "single_thread": {"value": one["value"], "cores": 1, "sample": one["sample"]}}
    if rank == 0:
        for k in [k for k in line if k.startswith("_")]:
            line.pop(k)
        print(json.dumps(line), flush=True)
    if world > 1:
        env.dist.barrier()
        env.dist.destroy_process_group()


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
    cores = oracle.set_threads(oracle.host_cores())
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
    method = args.method or "direct"
    regression.set_default_method(method)
    mine = nseries // world            # series are independent: partition, no collective
    x, y = make_batched(mine, n, first=rank * mine)
    ns = _lib.load().cf_lowess_span(n, f)
    stream = torch_stream_handle()
    b = regression.batched_lowess(device=local, stream=stream)
    b.set_method(method)
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
        line["config"]["method"] = method
        if method != "direct":
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
            cores = oracle.set_threads(oracle.host_cores())
            t0 = time.perf_counter()
            oracle.lowess_mt(x[0], y[0], f, nsteps)
            dt = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": n / dt, "unit": "points/s", "cores": cores,
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
