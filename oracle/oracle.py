"""ctypes binding of oracle/liblowess_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module.  The product package
(curvefit_b200) must never import it.  PARITY UNPINNED: see the header of
oracle/lowess_oracle.c (external anchor: the netlib lowess.f worked example).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblowess_oracle.so")

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build(force: bool = False) -> str:
    """Compile the C restatement with oracle/Makefile (gcc, no FMA contraction)."""
    srcs = [os.path.join(_HERE, f) for f in ("lowess_oracle.c", "stats_oracle.c")]
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(f) for f in srcs):
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, env=env,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        # idle OpenMP workers must sleep, not spin: spinning threads on a shared
        # host distort the timed CPU baseline
        os.environ.setdefault("OMP_WAIT_POLICY", "passive")
        L = C.CDLL(_SO)
        L.cf_oracle_span.argtypes = [C.c_int, C.c_double]
        L.cf_oracle_span.restype = C.c_int
        L.cf_oracle_lowess.argtypes = [_dp, _dp, C.c_int, C.c_double, C.c_int, C.c_double,
                                       _dp, _dp, _dp, _ip, _ip, _ip, _dp]
        L.cf_oracle_lowess.restype = C.c_int
        L.cf_oracle_lowess_mt.argtypes = [_dp, _dp, C.c_int, C.c_double, C.c_int,
                                          _dp, _dp, _dp, C.c_int, _dp]
        L.cf_oracle_lowess_mt.restype = C.c_int
        L.cf_oracle_lowest_at.argtypes = [_dp, _dp, C.c_int, C.c_double, _dp, _ip, C.c_int,
                                          _dp, _ip, _ip, _ip]
        L.cf_oracle_lowest_at.restype = C.c_int
        L.cf_oracle_sort_ind.argtypes = [_dp, _ip, C.c_int]
        L.cf_oracle_sort_ind.restype = None
        L.cf_oracle_ls_init.argtypes = [_dp, _dp, C.c_int, C.c_int, _dp, _dp]
        L.cf_oracle_ls_init.restype = C.c_int
        L.cf_oracle_ls_smooth.argtypes = [_dp, _dp, C.c_int, C.c_double, _dp, _dp, _dp]
        L.cf_oracle_ls_smooth.restype = C.c_int
        L.cf_oracle_moving_average.argtypes = [_dp, C.c_int, C.c_int]
        L.cf_oracle_moving_average.restype = C.c_int
        L.cf_oracle_median.argtypes = [_dp, C.c_int, C.c_int]
        L.cf_oracle_median.restype = C.c_double
        L.cf_oracle_set_threads.argtypes = [C.c_int]
        L.cf_oracle_set_threads.restype = C.c_int
        L.cf_oracle_robust_weights.argtypes = [_dp, _dp, C.c_int]
        L.cf_oracle_robust_weights.restype = C.c_double
        L.cf_oracle_max_threads.argtypes = []
        L.cf_oracle_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def span(n: int, f: float) -> int:
    return int(lib().cf_oracle_span(int(n), float(f)))


def lowess(x, y, f, nsteps=2, delta=0.0, want_windows=False):
    """Sequential reference lowess().  Returns dict(ys, rw, res, cmad[, nleft, nright, nrt])."""
    x = _f64(x)
    y = _f64(y)
    n = x.size
    ys = np.zeros(n)
    rw = np.zeros(n)
    res = np.zeros(n)
    cmad = np.zeros(max(nsteps, 1))
    if want_windows:
        nl = np.zeros(n, np.int32)
        nr = np.zeros(n, np.int32)
        nrt = np.zeros(n, np.int32)
        rc = lib().cf_oracle_lowess(_d(x), _d(y), n, f, nsteps, delta, _d(ys), _d(rw), _d(res),
                                    _i(nl), _i(nr), _i(nrt), _d(cmad))
    else:
        rc = lib().cf_oracle_lowess(_d(x), _d(y), n, f, nsteps, delta, _d(ys), _d(rw), _d(res),
                                    None, None, None, _d(cmad))
    assert rc == 0, rc
    out = dict(ys=ys, rw=rw, res=res, cmad=cmad[:nsteps])
    if want_windows:
        out.update(nleft=nl, nright=nr, nrt=nrt)
    return out


def lowess_mt(x, y, f, nsteps=2, nthreads=0):
    """All-cores variant (delta = 0), bit-identical to lowess() for ascending x."""
    x = _f64(x)
    y = _f64(y)
    n = x.size
    ys = np.zeros(n)
    rw = np.zeros(n)
    res = np.zeros(n)
    cmad = np.zeros(max(nsteps, 1))
    rc = lib().cf_oracle_lowess_mt(_d(x), _d(y), n, f, nsteps, _d(ys), _d(rw), _d(res),
                                   int(nthreads), _d(cmad))
    assert rc == 0, rc
    return dict(ys=ys, rw=rw, res=res, cmad=cmad[:nsteps])


def lowest_at(x, y, f, idx, rw=None):
    """Teacher-forced single pass at 0-based points idx, with weights rw (None = first pass)."""
    x = _f64(x)
    y = _f64(y)
    idx = np.ascontiguousarray(idx, dtype=np.int32)
    m = idx.size
    ys = np.zeros(m)
    nl = np.zeros(m, np.int32)
    nr = np.zeros(m, np.int32)
    nrt = np.zeros(m, np.int32)
    rwp = None
    if rw is not None:
        rw = _f64(rw)
        rwp = _d(rw)
    rc = lib().cf_oracle_lowest_at(_d(x), _d(y), x.size, f, rwp, _i(idx), m, _d(ys),
                                   _i(nl), _i(nr), _i(nrt))
    assert rc == 0, rc
    return dict(ys=ys, nleft=nl, nright=nr, nrt=nrt)


def sort_ind(x):
    """Reference sort(x, ind, .true.): returns (sorted copy, 1-based tracking indices)."""
    xs = _f64(x).copy()
    ind = np.arange(1, xs.size + 1, dtype=np.int32)
    lib().cf_oracle_sort_ind(_d(xs), _i(ind), xs.size)
    return xs, ind


def ls_init(x, y, srt=True):
    x = _f64(x)
    y = _f64(y)
    mx = np.zeros_like(x)
    my = np.zeros_like(y)
    rc = lib().cf_oracle_ls_init(_d(x), _d(y), x.size, int(bool(srt)), _d(mx), _d(my))
    assert rc == 0, rc
    return mx, my


def max_threads() -> int:
    return int(lib().cf_oracle_max_threads())


def set_threads(nthreads: int) -> int:
    """Threads of lowest_at / lowess_mt from now on (overrides OMP_NUM_THREADS, which
    torchrun exports as 1).  Returns the value in effect."""
    return int(lib().cf_oracle_set_threads(int(nthreads)))


def host_cores() -> int:
    """Cores this process may run on."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def robust_weights(res):
    """:697-713 on its own: (cmad, rw) from the residuals, single-threaded full sort."""
    res = _f64(res)
    rw = np.zeros_like(res)
    c = lib().cf_oracle_robust_weights(_d(res), _d(rw), res.size)
    return float(c), rw


def moving_average(x, npts):
    """moving_average_1 (src/curvefit_regression.f90:427-478) on a copy; returns
    (smoothed, flag) with flag = 0 or the reference's error code."""
    out = _f64(x).copy()
    rc = lib().cf_oracle_moving_average(_d(out), out.size, int(npts))
    return out, rc


def median(x, srt=True):
    """median_dbl (src/curvefit_statistics.f90:258-295); returns (z, x after the call)."""
    a = _f64(x).copy()
    z = lib().cf_oracle_median(_d(a), a.size, 1 if srt else 0)
    return float(z), a
