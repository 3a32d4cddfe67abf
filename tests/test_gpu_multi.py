"""Multi-GPU parity, skipped on boxes with fewer than two GPUs.

 * cf_lowess_smooth_multi -- one process, one host thread per device, peers through
   cudaDeviceEnablePeerAccess, halo-limited uploads -- against cf_lowess_smooth on one GPU
   (bitwise for distinct x) and against the oracle;
 * the one-process-per-GPU path (torchrun + CUDA IPC, both transports, both evaluators):
   scripts/check_sharded.py as a subprocess."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from curvefit_b200 import datagen
from tests import cases

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_ngpus() < 2, reason="needs at least two GPUs")


def _smooth(fn, x, y, f, nsteps, *extra):
    from curvefit_b200 import _lib
    n = x.size
    ys, rw, res = np.full(n, np.nan), np.full(n, np.nan), np.full(n, np.nan)
    p = lambda a: a.ctypes.data_as(_lib.c_dp)
    rc = fn(p(x), p(y), n, float(f), int(nsteps), *extra, p(ys), p(rw), p(res))
    assert rc == 0, _lib.last_error()
    return ys, rw, res


@needs2
@pytest.mark.parametrize("method", ["direct", "tree"])
@pytest.mark.parametrize("kind,n,f,nsteps", [("grid", 200_000, 0.05, 3), ("uniform", 150_001, 0.2, 2),
                                             ("clustered", 100_000, 0.1, 3), ("uniform", 60_000, 0.9, 1)])
def test_smooth_multi_equals_single_gpu(oracle, method, kind, n, f, nsteps):
    from curvefit_b200 import _lib
    lib = _lib.load()
    x = cases.make_x(kind, n, 7)
    y = cases.make_y(x, 7)
    lib.cf_set_default_method(_lib.CF_METHOD_TREE if method == "tree" else _lib.CF_METHOD_DIRECT)
    try:
        one = _smooth(lib.cf_lowess_smooth, x, y, f, nsteps)
        for ndev in sorted({2, _ngpus()}):
            got = _smooth(lib.cf_lowess_smooth_multi, x, y, f, nsteps, None, ndev)
            distinct = np.all(np.diff(x) > 0)
            for a, b, name in zip(got, one, ("ys", "rw", "res")):
                assert not np.isnan(a).any(), name
                ok, d = cases.close(a, b, 1e-12 if name != "rw" else 1e-9)
                assert ok, (name, ndev, d)
            if distinct and method == "tree":
                # per-point sums never cross devices and the tree's nodes do not depend on the
                # partition (the direct kernel's K-way window split does: tolerance above)
                assert np.array_equal(got[0], one[0])
        if n <= 100_000:
            ref = oracle.lowess(x, y, f, nsteps)
            ok, d = cases.close(got[0], ref["ys"])
            assert ok, d
    finally:
        lib.cf_set_default_method(_lib.CF_METHOD_DIRECT)
        lib.cf_lowess_multi_release()


@needs2
def test_smooth_multi_ties_and_device_lists(oracle):
    from curvefit_b200 import _lib
    lib = _lib.load()
    n, f, nsteps = 40_000, 0.15, 2
    x = cases.make_x("ties", n, 3)
    y = cases.make_y(x, 3)
    ref = oracle.lowess(x, y, f, nsteps)
    devs = (C.c_int32 * 2)(1, 0)                              # explicit list, reversed order
    ys, rw, res = _smooth(lib.cf_lowess_smooth_multi, x, y, f, nsteps, devs, 2)
    ok, d = cases.close(ys, ref["ys"])
    assert ok, d
    assert np.array_equal(res, y - ys)
    # repeated call on the cached handles with other data of the same length
    y2 = cases.make_y(x, 4)
    ys2, _, _ = _smooth(lib.cf_lowess_smooth_multi, x, y2, f, nsteps, devs, 2)
    ok, d = cases.close(ys2, oracle.lowess(x, y2, f, nsteps)["ys"])
    assert ok, d
    # bad lists
    p = lambda a: a.ctypes.data_as(_lib.c_dp)
    dup = (C.c_int32 * 2)(0, 0)
    out = np.empty(n)
    assert lib.cf_lowess_smooth_multi(p(x), p(y), n, f, nsteps, dup, 2, p(out), p(out), p(out)) == 201
    far = (C.c_int32 * 2)(0, 99)
    assert lib.cf_lowess_smooth_multi(p(x), p(y), n, f, nsteps, far, 2, p(out), p(out), p(out)) == 201
    # short series and one-device lists take the single-GPU path
    xs, ysm = x[:3000].copy(), y[:3000].copy()
    a = _smooth(lib.cf_lowess_smooth_multi, xs, ysm, 0.3, 2, None, 0)
    b = _smooth(lib.cf_lowess_smooth, xs, ysm, 0.3, 2)
    assert np.array_equal(a[0], b[0])
    lib.cf_lowess_multi_release()


@needs2
def test_smooth_multi_falls_back_to_one_gpu_when_a_device_is_full():
    """A GPU that another job has filled must not turn smooth() into an error: the sharded context
    cannot be built (cudaMalloc fails on device 1), the call runs on one GPU with the same results,
    and once the memory is free again the next call shards as before."""
    import torch
    from curvefit_b200 import _lib
    lib = _lib.load()
    n, f, nsteps = 4_000_000, 0.001, 2                         # ~240 MB of buffers per device
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x, 5)
    lib.cf_set_default_method(_lib.CF_METHOD_TREE)
    hog = None
    try:
        lib.cf_lowess_multi_release()
        one = _smooth(lib.cf_lowess_smooth, x, y, f, nsteps)
        torch.cuda.init()
        with torch.cuda.device(1):
            torch.cuda.empty_cache()
            free, _ = torch.cuda.mem_get_info()
            hog = torch.empty(free - (128 << 20), dtype=torch.uint8, device="cuda:1")
        devs = (C.c_int32 * 2)(0, 1)
        got = _smooth(lib.cf_lowess_smooth_multi, x, y, f, nsteps, devs, 2)
        assert lib.cf_lowess_multi_active() == 0               # no sharded context survived
        for a, b in zip(got, one):
            assert np.array_equal(a, b)
        del hog
        hog = None
        with torch.cuda.device(1):
            torch.cuda.empty_cache()
        again = _smooth(lib.cf_lowess_smooth_multi, x, y, f, nsteps, devs, 2)
        assert lib.cf_lowess_multi_active() == 2               # out of memory is not remembered
        assert np.array_equal(again[0], one[0])                # tree evaluator: partition invariant
    finally:
        del hog
        lib.cf_set_default_method(_lib.CF_METHOD_DIRECT)
        lib.cf_lowess_multi_release()


@needs2
def test_one_process_per_gpu_paths_match_single_gpu():
    """torchrun, 2 ranks: NCCL transport + direct, peer-memory transport + direct / tree, seven
    series each (three of them with delta > 0), every rank compares with its own single-GPU run."""
    env = dict(os.environ)
    env.pop("OMP_NUM_THREADS", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "scripts", "check_sharded.py")]
    r = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "sharded parity OK" in r.stdout
