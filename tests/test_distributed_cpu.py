"""CPU (gloo, world_size 2 and 3) test of the multi-GPU HOST logic in
curvefit_b200/distributed.py: the range partition, the radix-select protocol
over all-reduced histograms and the all-gather of robustness weights.  The
device stages are replaced by a numpy stand-in built on the oracle (tests may
use the oracle; the product never does), so the sharded result must equal the
sequential oracle bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from curvefit_b200.distributed import ShardedLowess, partition
from tests import cases

# the select geometry of csrc/common.cuh (one series): keys of 63 informative bits, round 0
# takes the top 15 of every residual, rounds 1..4 take 12 bits each of the candidates only
TOP, B0, W = 63, 15, 12
ROUNDS = 1 + (TOP - B0 + W - 1) // W


class NumpyStages:
    """Same attributes/methods as DeviceStages, computed with numpy + the oracle."""

    def __init__(self, x, y, rank, nranks):
        from oracle import oracle
        self.o = oracle
        self.x, self.y = x, y
        self.n = x.size
        self.rank, self.nranks = rank, nranks
        self.block, parts = partition(self.n, nranks)
        self.begin, self.end = parts[rank]
        self.hist0 = torch.zeros(1 << B0, dtype=torch.int32)
        self.hist = torch.zeros(2 << W, dtype=torch.int32)
        self.rw = torch.zeros(nranks * self.block, dtype=torch.float64)
        self.ys = torch.zeros(self.n, dtype=torch.float64)
        self.res = torch.zeros(self.n, dtype=torch.float64)
        self.cmad = torch.zeros(3, dtype=torch.float64)
        self.rounds = ROUNDS
        self.prefix = [0, 0]
        self.krank = [0, 0]
        self.cand = np.zeros(0, np.uint64)

    def begin_run(self, f):
        self.f = f

    def fit_pass(self, use_rw):
        idx = np.arange(self.begin, self.end, dtype=np.int32)
        rw = self.rw.numpy()[: self.n].copy() if use_rw else None
        if idx.size:
            t = self.o.lowest_at(self.x, self.y, self.f, idx, rw)
            self.ys.numpy()[self.begin:self.end] = t["ys"]
            self.res.numpy()[self.begin:self.end] = self.y[self.begin:self.end] - t["ys"]

    def hist_view(self, r):
        return self.hist0 if r == 0 else self.hist

    @staticmethod
    def _geom(k):
        """(shift, width) of round k >= 1."""
        rounds = ROUNDS - 1
        width = (TOP - B0) - W * (rounds - 1) if k == 1 else W
        return W * (rounds - k), width

    def hist_round(self, r):
        if r == 0:
            keys = np.abs(self.res.numpy()[self.begin:self.end]).view(np.uint64)
            self.hist0.numpy()[:] = np.bincount((keys >> np.uint64(TOP - B0)).astype(np.int64), minlength=1 << B0)
            return
        shift, width = self._geom(r)
        hi = self.cand >> np.uint64(shift + width)
        d = ((self.cand >> np.uint64(shift)) & np.uint64((1 << width) - 1)).astype(np.int64)
        h = self.hist.numpy()
        h[:] = 0
        same = self.prefix[0] == self.prefix[1]
        for t in range(1 if same else 2):
            h[t << W:(t + 1) << W] = np.bincount(d[hi == np.uint64(self.prefix[t])], minlength=1 << W)

    def pick_round(self, r):
        if r == 0:
            c = np.cumsum(self.hist0.numpy().astype(np.int64))
            for t, k in enumerate([(self.n - 1) // 2, self.n // 2]):
                b = int(np.searchsorted(c, k, side="right"))
                self.krank[t] = k - (int(c[b - 1]) if b > 0 else 0)
                self.prefix[t] = b
            keys = np.abs(self.res.numpy()[self.begin:self.end]).view(np.uint64)
            top = keys >> np.uint64(TOP - B0)
            self.cand = keys[(top == np.uint64(self.prefix[0])) | (top == np.uint64(self.prefix[1]))]
            return
        _, width = self._geom(r)
        h = self.hist.numpy().astype(np.int64)
        same = self.prefix[0] == self.prefix[1]
        for t in range(2):
            src = h[:1 << W] if (same or t == 0) else h[1 << W:]
            c = np.cumsum(src)
            b = int(np.searchsorted(c, self.krank[t], side="right"))
            self.krank[t] -= int(c[b - 1]) if b > 0 else 0
            self.prefix[t] = (self.prefix[t] << width) | b
        if r == ROUNDS - 1:
            lo, hi = (np.array(self.prefix, dtype=np.uint64).view(np.float64))
            cmad = 3.0 * (hi + lo)
            self.cmad[:] = torch.tensor([cmad, 0.001 * cmad, 0.999 * cmad])

    def weights(self):
        cmad, c1, c9 = self.cmad.tolist()
        r = np.abs(self.res.numpy()[self.begin:self.end])
        with np.errstate(invalid="ignore", divide="ignore"):
            t = r / cmad
            u = 1.0 - t * t
            w = np.where(r <= c1, 1.0, np.where(r > c9, 0.0, u * u))
        self.rw.numpy()[self.begin:self.end] = w

    def sync(self):
        pass

    # host-buffer path (smooth_host with shared_out): only the span is "uploaded", only the block read back
    def upload_span(self, x, y, f):
        ns = max(min(int(f * float(self.n)), self.n), 2)
        lo, hi = max(self.begin - ns, 0), min(self.end - 1 + ns, self.n - 1)
        lo = int(np.searchsorted(x[:lo], x[lo], side="left")) if lo > 0 else 0
        hi = hi + int(np.searchsorted(x[hi:], x[hi], side="right")) - 1
        self.x = np.where((np.arange(self.n) < lo), -np.inf, np.where(np.arange(self.n) > hi, np.inf, x))
        self.x[0], self.x[-1] = x[0], x[-1]
        self.y = np.where((np.arange(self.n) >= lo) & (np.arange(self.n) <= hi), y, np.nan)
        return hi - lo + 1

    def fetch_block(self, ys, rw, res):
        b, e = self.begin, self.end
        ys[b:e] = self.ys.numpy()[b:e]
        rw[b:e] = self.rw.numpy()[b:e]
        res[b:e] = self.res.numpy()[b:e]


def _worker(rank, world, port, n, f, nsteps, kind, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x = cases.make_x(kind, n, 21)
        y = cases.make_y(x, 21)
        job = ShardedLowess(NumpyStages(x, y, rank, world))
        ys, res = job.run(f, nsteps)
        rw = job.s.rw[:n].clone()
        q.put((rank, ys.numpy().copy(), res.numpy().copy(), rw.numpy().copy(), job.collectives,
               job.s.cmad[0].item()))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,n,f,nsteps,kind", [(2, 1500, 0.2, 2, "uniform"),
                                                   (3, 2000, 0.1, 3, "grid"),
                                                   (2, 700, 0.3, 2, "ties")])
def test_sharded_lowess_over_gloo_equals_sequential_oracle(oracle, world, n, f, nsteps, kind):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, f, nsteps, kind, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x = cases.make_x(kind, n, 21)
    y = cases.make_y(x, 21)
    ref = oracle.lowess(x, y, f, nsteps)
    for rank, ys, res, rw, ncoll, cmad in outs:
        assert np.array_equal(ys, ref["ys"]), rank
        assert np.array_equal(res, ref["res"]), rank
        assert np.array_equal(rw, ref["rw"]), rank
        assert cmad == ref["cmad"][-1]
        # per iteration: ROUNDS histogram all-reduces + 1 all-gather; 2 result gathers
        assert ncoll == nsteps * (ROUNDS + 1) + 2


def test_partition_is_tile_aligned_and_covers_everything():
    for n, g in [(1_000_000, 8), (100, 2), (513, 3), (100_000_000, 8), (5, 4)]:
        block, parts = partition(n, g)
        assert block % 512 == 0 and len(parts) == g
        assert parts[0][0] == 0 and parts[-1][1] == n
        for (b0, e0), (b1, e1) in zip(parts, parts[1:]):
            assert e0 == b1 and b0 <= e0
        assert all(b % 512 == 0 or b == n for b, _ in parts)


def _worker_host(rank, world, port, n, f, nsteps, q):
    from curvefit_b200.distributed import SharedHostArray
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        x = cases.make_x("uniform", n, 5)
        y = cases.make_y(x, 5)
        outs = [SharedHostArray(n) for _ in range(3)]
        for o in outs:
            if rank == 0:
                o.array[:] = np.nan
        dist.barrier()
        job = ShardedLowess(NumpyStages(x.copy(), y.copy(), rank, world))
        job.smooth_host(x, y, f, nsteps, tuple(o.array for o in outs), shared_out=True)
        # every rank sees the whole result although each wrote only its block
        q.put((rank, [o.array.copy() for o in outs], job.last_upload))
        for o in outs:
            o.close()
    finally:
        dist.destroy_process_group()


def test_smooth_host_shared_output_uploads_halo_only_and_assembles_blocks(oracle):
    """world_size 2 over gloo: each rank receives only its block + halo (the rest of x is
    +-inf, of y NaN -- the oracle stand-in would propagate them if a window reached there)
    and writes only its own block of the shared result."""
    world, n, f, nsteps = 2, 3000, 0.1, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_host, args=(r, world, port, n, f, nsteps, q)) for r in range(world)]
    for p in procs:
        p.start()
    outs = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x = cases.make_x("uniform", n, 5)
    y = cases.make_y(x, 5)
    ref = oracle.lowess(x, y, f, nsteps)
    for rank, (ys, rw, res), uploaded in outs:
        assert np.array_equal(ys, ref["ys"]), rank
        assert np.array_equal(res, ref["res"]), rank
        assert np.array_equal(rw, ref["rw"]), rank
        assert uploaded < n                   # halo-limited, not the whole series
