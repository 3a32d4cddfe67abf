"""Multi-GPU parity check, run under torchrun (one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 scripts/check_sharded.py
Every rank runs the range-sharded smooth (NCCL collectives between the device stages)
and compares it with the same series smoothed on its own GPU alone and with the oracle
at sampled points."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from curvefit_b200 import datagen, lowess_smoothing
from curvefit_b200.distributed import DeviceStages, ShardedLowess, torch_stream_handle


def main():
    rank = int(os.environ["RANK"])
    world = int(os.environ["WORLD_SIZE"])
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    # (n, f, nsteps, x family, delta as a fraction of the x range; 0 = every point fitted)
    cases = [(50_000, 0.1, 3, "random", 0.0), (200_000, 0.05, 2, "grid", 0.0), (3000, 0.3, 2, "ties", 0.0),
             (300_000, 0.2, 3, "random", 0.0),
             (200_000, 0.1, 2, "random", 0.002), (60_000, 0.3, 3, "ties", 0.01), (100_000, 0.05, 2, "grid", 0.0005)]
    for transport, method in [("nccl", "direct"), ("p2p", "direct"), ("p2p", "tree")]:
      for n, f, nsteps, kind, dfrac in cases:
          if kind == "grid":
              x = datagen.grid_x(n)
          elif kind == "ties":
              x = np.sort(np.round(datagen.random_x(n, 5) * 400) / 400)
          else:
              x = np.sort(datagen.random_x(n, 3))
          y = datagen.noisy_y(x, 3)
          stream = torch_stream_handle()
          solo = lowess_smoothing(device=local)
          solo.set_method(method)
          solo.initialize(x, y, srt=False)
          delta = dfrac * float(x[-1] - x[0])
          if delta > 0:
              solo.set_delta(delta)
          ys_ref = solo.smooth(f, nsteps=nsteps)
          rw_ref, res_ref = solo.m_weights.copy(), solo.m_residuals.copy()

          fit = lowess_smoothing(device=local, stream=stream)
          fit.set_method(method)
          fit.initialize(x, y, srt=False)
          if delta > 0:
              fit.set_delta(delta)      # every rank derives the whole fitted-point list, fits its block's slots
          stages = DeviceStages(fit, rank, world)
          if transport == "p2p":
              assert stages.attach_peers(), "peer attach failed"
          job = ShardedLowess(stages)
          ys, res = job.run(f, nsteps, with_rw=True)
          torch.cuda.synchronize()
          ys, res = ys.cpu().numpy(), res.cpu().numpy()
          rw = job.s.rw[:n].cpu().numpy()
          cmad = job.s.cmad.cpu().numpy()[0]
          scale = np.max(np.abs(ys_ref))
          d_ys = np.max(np.abs(ys - ys_ref)) / scale
          d_rw = np.max(np.abs(rw - rw_ref))
          d_res = np.max(np.abs(res - res_ref)) / scale
          d_c = abs(cmad - solo.last_stats.cmad_last) / solo.last_stats.cmad_last
          perr = stages.peer_error() if transport == "p2p" else False
          good = d_ys <= 1e-12 and d_rw <= 1e-9 and d_res <= 1e-12 and d_c <= 1e-12 and not perr
          ok &= bool(good)
          print(f"[rank {rank}/{world}] {transport}/{method} n={n} f={f} nsteps={nsteps} {kind} delta={delta:.3g}: |dys|={d_ys:.2e} |drw|={d_rw:.2e} "
                f"|dres|={d_res:.2e} dcmad={d_c:.2e} bitwise_ys={np.array_equal(ys, ys_ref)} "
                f"collectives={job.collectives} {'OK' if good else 'FAIL'}", flush=True)
          stages.detach_peers()     # collective; before any rank frees its buffers
          del job, stages, fit, solo
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if flag.item() != 1:
        sys.exit(1)
    if rank == 0:
        print("sharded parity OK")


if __name__ == "__main__":
    main()
