"""One series range-sharded over the GPUs of one box.

One process per GPU (torchrun); every rank holds the whole sorted x, y (16 B
per point) and the whole robustness-weight vector, fits only its own block
[begin, end) of output points, and per robustness iteration exchanges
  * what the exact select of the two middle order statistics of |res| needs
    (src/curvefit_regression.f90:697-703 -- the reference's full sort): the
    digit histograms of the radix select (round 0: 2^15 counters over every
    residual; later rounds: 2 x 2^12 counters over the few candidates), summed
    over the ranks -> every rank resolves the same cmad;
  * its block of new robustness weights (all-gather) -- the superset of the
    (ns-1)-point halo the neighbours need for :551.
Two transports: "p2p" (default on GPUs) -- the kernels do the exchanges
themselves through NVLink peer memory inside ONE cf_lowess_run call per rank
(csrc/peer.cu: round-0 histogram all-reduce fused into the residual pass's scan,
candidate keys and halo weights stored straight into every rank's memory, device
barriers; torch.distributed only carries the CUDA IPC handles once) -- and "nccl"
(torch.distributed collectives between the device stages -- the baseline, and
what the gloo CPU tests exercise).  The device stages are libcurvefit_cuda's cf_lowess_stage_* entry points.  Per-point
sums never cross ranks, so ys is bit-identical for any number of ranks.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch
import torch.distributed as dist

from . import _lib


def partition(n: int, nranks: int, tile: int = 512):
    """Equal blocks of output points, rounded up to the fit kernel's tile so that
    tile boundaries (and therefore rounding) do not depend on nranks.  Mirrors
    compute_shard() in csrc/api.cu.  Returns (block, [(begin, end), ...])."""
    per = -(-n // nranks)
    per = max(-(-per // tile) * tile, tile)
    out = []
    for r in range(nranks):
        b = min(r * per, n)
        out.append((b, min(b + per, n)))
    return per, out


def torch_stream_handle() -> int:
    """cudaStream_t of torch's current stream in the form libcurvefit_cuda expects.
    torch's default stream is the legacy default stream, whose handle is 0 -- which
    the C ABI reads as "create a private stream".  cudaStreamLegacy (0x1) names the
    same stream explicitly, so kernels and NCCL collectives stay ordered."""
    h = int(torch.cuda.current_stream().cuda_stream)
    return h if h != 0 else 1


class _DevArray:
    """Zero-copy view of a raw device pointer for torch.as_tensor."""

    def __init__(self, ptr: int, count: int, typestr: str):
        self.__cuda_array_interface__ = {
            "shape": (count,), "typestr": typestr, "data": (int(ptr), False), "version": 3,
            "strides": None,
        }


class DeviceStages:
    """The CUDA stages of one rank, bound to a lowess_smoothing handle that was
    created on torch's current stream."""

    def __init__(self, fit, rank: int, nranks: int):
        self.lib = _lib.load()
        self.fit = fit
        self.h = fit._h
        self.n = fit.get_count()
        self.rank, self.nranks = rank, nranks
        self._check(self.lib.cf_lowess_set_shard(self.h, rank, nranks))
        b, e, blk = C.c_int32(), C.c_int32(), C.c_int32()
        self.lib.cf_lowess_shard(self.h, C.byref(b), C.byref(e), C.byref(blk))
        self.begin, self.end, self.block = b.value, e.value, blk.value
        dev = torch.device("cuda", torch.cuda.current_device())
        buf = lambda which: self.lib.cf_lowess_buffer(self.h, which)
        self.hist0 = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_HIST0), self.lib.cf_lowess_hist0_len(), "<i4"), device=dev)
        self.hist = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_HIST), self.lib.cf_lowess_hist_len(), "<i4"), device=dev)
        self.rw = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_RW), nranks * self.block, "<f8"), device=dev)
        self.ys = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_YS), self.n, "<f8"), device=dev)
        self.res = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_RES), self.n, "<f8"), device=dev)
        self.cmad = torch.as_tensor(_DevArray(buf(_lib.CF_BUF_CMAD), 3, "<f8"), device=dev)
        self.rounds = self.lib.cf_lowess_select_rounds()
        self.launches = 0
        self.p2p = False

    def attach_peers(self, group: Optional[dist.ProcessGroup] = None) -> bool:
        """Map every rank's exchange block, rw, ys and res into this process (CUDA IPC)
        so that the stages exchange through peer memory.  Collective.  Returns False
        (and leaves the NCCL transport in place) if any rank could not attach."""
        if self.nranks == 1:
            return False
        nb = self.lib.cf_lowess_peer_handle_bytes()
        mine = (C.c_uint8 * nb)()
        rc = self.lib.cf_lowess_peer_export(self.h, mine)
        dev = torch.device("cuda", torch.cuda.current_device())
        own = torch.frombuffer(bytearray(bytes(mine)), dtype=torch.uint8).to(dev)
        allh = torch.empty(nb * self.nranks, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allh, own, group=group)
        if rc == 0:
            blob = bytes(allh.cpu().numpy().tobytes())
            rc = self.lib.cf_lowess_peer_attach(self.h, blob, self.nranks)
        ok = torch.tensor([1 if rc == 0 else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)      # also the post-attach barrier
        torch.cuda.synchronize()
        self.p2p = bool(ok.item() == 1)
        if not self.p2p:
            self.lib.cf_lowess_peer_detach(self.h)
        return self.p2p

    def detach_peers(self, group: Optional[dist.ProcessGroup] = None) -> None:
        """Collective: no rank unmaps while a peer may still store, and no rank frees
        its buffers (re-init, close) while a peer still has them mapped."""
        if not self.p2p:
            return
        torch.cuda.synchronize()
        dist.barrier(group=group)
        self.lib.cf_lowess_peer_detach(self.h)
        self.p2p = False
        dist.barrier(group=group)

    def upload(self, x: np.ndarray, y: np.ndarray) -> None:
        """ls_init with srt = .false. on the existing handle (buffers stay in place when n
        is unchanged, so peer mappings and the tensor views above remain valid)."""
        assert x.size == self.n and y.size == self.n, "series length is fixed for a sharded handle"
        self._check(self.lib.cf_lowess_init(self.h, x.ctypes.data_as(_lib.c_dp), y.ctypes.data_as(_lib.c_dp),
                                            self.n, 0))

    def upload_span(self, x: np.ndarray, y: np.ndarray, f: float) -> int:
        """ls_init with srt = .false., uploading only the elements this rank's windows reach
        (block + the (ns-1)-point halo either side, cf_lowess_shard_span); returns how many."""
        assert x.size == self.n and y.size == self.n, "series length is fixed for a sharded handle"
        first, last = C.c_int32(), C.c_int32()
        xp, yp = x.ctypes.data_as(_lib.c_dp), y.ctypes.data_as(_lib.c_dp)
        self._check(self.lib.cf_lowess_shard_span(self.h, xp, self.n, float(f), C.byref(first), C.byref(last)))
        self._check(self.lib.cf_lowess_init_span(self.h, xp, yp, self.n, first.value, last.value))
        return last.value - first.value + 1

    def set_early_weights(self, rw) -> None:
        """The next fused run copies this rank's block of the final robustness weights into the
        full-length host array `rw` while the last fit pass runs (fetch_block completes it)."""
        self._check(self.lib.cf_lowess_set_early_weights(self.h, None if rw is None else rw.ctypes.data_as(_lib.c_dp)))

    def fetch_block(self, ys, rw, res) -> None:
        """Own block [begin, end) of ys, rw, res -> the same positions of full-length host arrays."""
        p = lambda a: None if a is None else a.ctypes.data_as(_lib.c_dp)
        self._check(self.lib.cf_lowess_fetch_block(self.h, p(ys), p(rw), p(res)))

    def fetch(self, ys, rw, res) -> None:
        p = lambda a: None if a is None else a.ctypes.data_as(_lib.c_dp)
        self._check(self.lib.cf_lowess_fetch(self.h, p(ys), p(rw), p(res)))

    def gather(self, what: int = _lib.CF_GATHER_YS | _lib.CF_GATHER_RES) -> None:
        self._check(self.lib.cf_lowess_stage_gather(self.h, what))

    def peer_error(self) -> bool:
        flag = C.c_int32(0)
        self._check(self.lib.cf_lowess_peer_error(self.h, C.byref(flag)))
        return flag.value != 0

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise RuntimeError(f"libcurvefit_cuda stage failed ({rc}): {_lib.last_error()}")

    def begin_run(self, f: float) -> None:
        self._check(self.lib.cf_lowess_stage_begin(self.h, float(f)))

    def fit_pass(self, use_rw: bool) -> None:
        self._check(self.lib.cf_lowess_stage_fit(self.h, 1 if use_rw else 0))

    def hist_round(self, r: int) -> None:
        self._check(self.lib.cf_lowess_stage_hist(self.h, r))

    def pick_round(self, r: int) -> None:
        self._check(self.lib.cf_lowess_stage_pick(self.h, r))

    def hist_view(self, r: int):
        """The counters round r's collective must sum over the ranks."""
        return self.hist0 if r == 0 else self.hist

    def weights(self) -> None:
        self._check(self.lib.cf_lowess_stage_weights(self.h))

    def run_fused(self, f: float, nsteps: int):
        """p2p transport: the whole smooth in one collective C call (cf_lowess_run on an
        attached handle); returns its cf_stats, raises when a peer wait timed out."""
        st = _lib.cf_stats()
        self._check(self.lib.cf_lowess_run(self.h, float(f), int(nsteps), C.byref(st)))
        return st

    def sync(self) -> None:
        self._check(self.lib.cf_lowess_sync(self.h))


class ShardedLowess:
    """lowess() over `stages.nranks` ranks.  `stages` is DeviceStages on GPUs; the
    CPU tests pass a stand-in with the same attributes to exercise this host
    logic over gloo."""

    def __init__(self, stages, group: Optional[dist.ProcessGroup] = None):
        self.s = stages
        self.group = group
        self.collectives = 0

    def run(self, f: float, nsteps: int = 2, gather: bool = True, with_rw: bool = False):
        s = self.s
        self.last_stats = None
        if getattr(s, "p2p", False):
            # the kernels exchange; a timed-out peer wait makes cf_lowess_run fail on the rank
            # that saw it -- the other ranks learn it from check_all()
            self.last_stats = s.run_fused(f, nsteps)
            return self.gather_results(with_rw) if gather else None
        multi = s.nranks > 1
        s.begin_run(f)
        for p in range(nsteps + 1):                           # :657
            s.fit_pass(p > 0)                                 # :662-695 on the block
            if p == nsteps:                                   # :696
                break
            for r in range(s.rounds):                         # :697-703 as a radix select
                s.hist_round(r)
                if multi:
                    dist.all_reduce(s.hist_view(r), op=dist.ReduceOp.SUM, group=self.group)
                    self.collectives += 1
                s.pick_round(r)
            s.weights()                                       # :704-713 on the block
            if multi:
                own = s.rw[s.rank * s.block:(s.rank + 1) * s.block]
                dist.all_gather_into_tensor(s.rw, own, group=self.group)
                self.collectives += 1
        if not gather:
            return None
        return self.gather_results(with_rw)

    def check_all(self) -> None:
        """Collective: raises on EVERY rank when any rank's device-side peer wait timed out
        (its results would be built on missing histograms or halo weights)."""
        s = self.s
        if not getattr(s, "p2p", False):
            return
        bad = torch.tensor([1 if s.peer_error() else 0], device=s.ys.device)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX, group=self.group)
        if int(bad.item()) != 0:
            raise RuntimeError("a peer-memory wait timed out on some rank: results are invalid")

    def smooth_host(self, x: np.ndarray, y: np.ndarray, f: float, nsteps: int = 2, out=None,
                    shared_out: bool = False):
        """smooth() with HOST buffers on every call.
        shared_out = False: each rank uploads the whole (sorted) series, the ranks smooth it
        together, and the ranks that pass `out` = (ys, rw, res) read the whole result back.
        shared_out = True: `out` arrays live in host memory ALL ranks map (SharedHostArray):
        each rank uploads only the elements its windows reach and writes only its own block of
        the result; a barrier ends the call, after which every rank sees the whole result.
        The series length must not change while peers are attached."""
        s = self.s
        if shared_out:
            self.last_upload = s.upload_span(x, y, f)
            ys, rw, res = out
            early = getattr(s, "p2p", False) and nsteps > 0 and rw is not None
            if early:
                s.set_early_weights(rw)            # leaves beside the last fit pass
            self.run(f, nsteps, gather=False)
            s.fetch_block(ys, None if early else rw, res)
            if s.nranks > 1:
                dist.barrier(group=self.group)
            return out
        s.upload(x, y)
        self.last_upload = s.n
        self.run(f, nsteps, gather=True, with_rw=nsteps > 0)
        if out is not None:
            s.fetch(*out)
        else:
            s.sync()
        return out

    def gather_results(self, with_rw: bool = False):
        """Full-length ys and res on every rank (block-padded all-gather).  with_rw: the
        p2p transport delivers only halo weights during the run, so the full-length
        m_weights vector is completed here when the caller wants it."""
        s = self.s
        if s.nranks == 1:
            return s.ys[: s.n], s.res[: s.n]
        if getattr(s, "p2p", False):
            # own block stored into every rank's ys / res (/ rw) + device barrier
            s.gather(_lib.CF_GATHER_YS | _lib.CF_GATHER_RES | (_lib.CF_GATHER_RW if with_rw else 0))
            return s.ys[: s.n], s.res[: s.n]
        outs = []
        for src in (s.ys, s.res):
            own = torch.zeros(s.block, dtype=src.dtype, device=src.device)
            cnt = s.end - s.begin
            if cnt > 0:
                own[:cnt] = src[s.begin:s.end]
            full = torch.empty(s.nranks * s.block, dtype=src.dtype, device=src.device)
            dist.all_gather_into_tensor(full, own, group=self.group)
            self.collectives += 1
            outs.append(full[: s.n])
        return outs[0], outs[1]


class SharedHostArray:
    """A float64 host array in POSIX shared memory that every rank of the job maps and
    page-locks (cudaHostRegister), so that each rank's device can copy its own block of a
    result straight into it.  Collective: rank 0 creates, the others attach."""

    def __init__(self, n: int, group: Optional[dist.ProcessGroup] = None):
        from multiprocessing import shared_memory
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        name = [None]
        self._shm = None
        if rank == 0:
            self._shm = shared_memory.SharedMemory(create=True, size=max(8 * n, 8))
            name[0] = self._shm.name
        if dist.is_initialized():
            dist.broadcast_object_list(name, src=0, group=group)
        if rank != 0:
            self._shm = shared_memory.SharedMemory(name=name[0])
            try:        # only the creator unlinks; keep this process's tracker from trying at exit
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:
                pass
        self.array = np.ndarray((n,), dtype=np.float64, buffer=self._shm.buf)
        self._owner = rank == 0
        self._registered = False
        if torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self.array.ctypes.data, 8 * n, 0)
            self._registered = int(rc) == 0
        if dist.is_initialized():
            dist.barrier(group=group)

    def close(self, group: Optional[dist.ProcessGroup] = None) -> None:
        if self._shm is None:
            return
        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.array.ctypes.data)
        self.array = None
        if dist.is_initialized():
            dist.barrier(group=group)
        self._shm.close()
        if self._owner:
            self._shm.unlink()
        self._shm = None
