"""Host-side mirror of the reference's `lowess_smoothing` type
(src/curvefit_regression.f90:218-247): same procedure names, argument meaning
and error behaviour; the numerical work runs in libcurvefit_cuda on a B200.

    fit = lowess_smoothing()
    fit.initialize(x, y)            # ls_init,  :740-811
    ys = fit.smooth(0.2)            # ls_smooth :832-868 (nsteps = 2, delta = 0)
    fit.get_count(); fit.get_x(i); fit.get_y(i); fit.get_residuals(buf)

Indices taken by get_x / get_y are 1-based like the reference's.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from .core import (CF_ARRAY_SIZE_ERROR, CF_DEVICE_ERROR, CF_INVALID_INPUT_ERROR,
                   CF_NO_DATA_DEFINED_ERROR, CF_OUT_OF_MEMORY_ERROR)
from .errors import errors

_MESSAGES = {
    CF_ARRAY_SIZE_ERROR: "Input array sizes must match.",          # :771
    CF_OUT_OF_MEMORY_ERROR: "Insufficient memory available.",      # :787, :863
    CF_NO_DATA_DEFINED_ERROR: "No data has been defined.",         # :853
}


_METHODS = {"direct": _lib.CF_METHOD_DIRECT, "tree": _lib.CF_METHOD_TREE, "auto": _lib.CF_METHOD_AUTO}


def _dp(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_lib.c_dp)


def _ip(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(_lib.c_ip)


def set_default_method(method) -> None:
    """Evaluator of handles created from now on and of the stateless calls (process-wide)."""
    code = _METHODS[method] if isinstance(method, str) else int(method)
    rc = _lib.load().cf_set_default_method(code)
    if rc != 0:
        raise RuntimeError(f"cf_set_default_method failed ({rc}): {_lib.last_error()}")


def moving_average(x: np.ndarray, npts: int, err: Optional[errors] = None) -> None:
    """moving_average_1 (src/curvefit_regression.f90:427-478): smooths the signal x IN
    PLACE with an npts-sample moving average (zeros before the start of the signal,
    running-update mean).  x must be a contiguous float64 array.  Errors as the reference:
    npts < 2 or npts > size(x) -> CF_INVALID_INPUT_ERROR through `err`, x untouched."""
    errmgr = err if err is not None else errors()             # :443-448
    if not (isinstance(x, np.ndarray) and x.dtype == np.float64 and x.flags.c_contiguous and x.ndim == 1):
        raise TypeError("x must be a contiguous 1-d float64 array (it is smoothed in place)")
    n = x.size
    if npts < 2 or npts > n:                                  # :452-457
        errmgr.report_error("moving_average_1", "The averaging size window must be at least 2, and less "
                            "than the size of the data set.", CF_INVALID_INPUT_ERROR)
        return
    rc = _lib.load().cf_moving_average(_dp(x), n, int(npts))
    if rc != 0:
        msg = _MESSAGES.get(rc) or _lib.last_error() or "CUDA device failure."
        errmgr.report_error("moving_average_1", msg, rc)


class lowess_smoothing:
    """Robust locally weighted scatterplot smoothing (type lowess_smoothing, :218-247)."""

    def __init__(self, device: int = -1, stream: int = 0):
        self._lib = _lib.load()            # ImportError if the CUDA library is not built
        self._device = device
        self._stream = stream
        self._h = C.c_void_p()             # created lazily: needs a GPU
        self.m_x: Optional[np.ndarray] = None          # :220
        self.m_y: Optional[np.ndarray] = None          # :223
        self.m_weights: Optional[np.ndarray] = None    # :225
        self.m_residuals: Optional[np.ndarray] = None  # :227
        self.m_delta = 0.0                              # :230
        self.m_init = False                             # :232
        self.last_stats: Optional[_lib.cf_stats] = None

    # -- plumbing -------------------------------------------------------------
    def _report(self, errmgr: errors, fcn: str, code: int) -> None:
        msg = _MESSAGES.get(code)
        if msg is None:
            msg = _lib.last_error() or "CUDA device failure."
        errmgr.report_error(fcn, msg, code)

    def _handle(self) -> int:
        if not self._h:
            rc = self._lib.cf_lowess_create(C.byref(self._h), self._device,
                                            C.c_void_p(self._stream) if self._stream else None)
            if rc != 0:
                self._h = C.c_void_p()
                return rc
        return 0

    def close(self) -> None:
        if self._h:
            self._lib.cf_lowess_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_method(self, method) -> None:
        """Evaluator of a lowest() pass: "direct" (O(n ns) sweep), "tree" (O(n log ns)
        moment tree, equal to rounding) or "auto".  Extension; see curvefit_cuda.h."""
        code = _METHODS[method] if isinstance(method, str) else int(method)
        rc = self._handle()
        if rc == 0:
            rc = self._lib.cf_lowess_set_method(self._h, code)
        if rc != 0:
            raise RuntimeError(f"cf_lowess_set_method failed ({rc}): {_lib.last_error()}")

    def set_delta(self, delta: float) -> None:
        """m_delta (:230): lowess's interpolation distance.  The reference fixes it at 0
        (:807; the setter is a stub, :808-809); delta > 0 fits only points more than delta
        apart and interpolates the rest (:674-690).  Extension; see curvefit_cuda.h."""
        rc = self._handle()
        if rc == 0:
            rc = self._lib.cf_lowess_set_delta(self._h, float(delta))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_set_delta failed ({rc}): {_lib.last_error()}")
        self.m_delta = float(delta)

    def fitted_points(self) -> np.ndarray:
        """1-based indices of the points lowest() ran at in the last delta > 0 smooth."""
        cnt = C.c_int32(0)
        rc = self._lib.cf_lowess_fitted(self._h, C.byref(cnt), None, 0)
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fitted failed ({rc}): {_lib.last_error()}")
        idx = np.zeros(cnt.value, np.int32)
        rc = self._lib.cf_lowess_fitted(self._h, C.byref(cnt), _ip(idx), idx.size)
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fitted failed ({rc}): {_lib.last_error()}")
        return idx

    # -- ls_init (:740-811) ---------------------------------------------------------
    def initialize(self, x, y, srt: bool = True, err: Optional[errors] = None) -> None:
        self.m_init = False                                   # :758
        errmgr = err if err is not None else errors()         # :762-766
        x = np.asarray(x, dtype=np.float64)
        y = np.asarray(y, dtype=np.float64)
        n = x.size
        if y.size != n or x.ndim != 1 or y.ndim != 1:         # :769-773
            self._report(errmgr, "ls_init", CF_ARRAY_SIZE_ERROR)
            return
        rc = self._handle()
        if rc == 0:
            xc = np.ascontiguousarray(x)
            yc = np.ascontiguousarray(y)
            rc = self._lib.cf_lowess_init(self._h, _dp(xc), _dp(yc), n, 1 if srt else 0)
        if rc == 0:
            # host mirrors serve get_x / get_y (:893-920) without a device call
            self.m_x = np.empty(n)
            self.m_y = np.empty(n)
            rc = self._lib.cf_lowess_fetch_xy(self._h, _dp(self.m_x), _dp(self.m_y))
        if rc != 0:
            self._report(errmgr, "ls_init", rc)
            return
        # allocated but not initialised in the reference (:780-784)
        self.m_weights = np.zeros(n)
        self.m_residuals = np.zeros(n)
        if self.m_delta != 0.0:                               # :807 -- initialize resets delta
            self._lib.cf_lowess_set_delta(self._h, 0.0)
        self.m_delta = 0.0
        self.m_init = True                                    # :810

    # -- ls_smooth (:832-868) -----------------------------------------------------
    def smooth(self, f: float, err: Optional[errors] = None, nsteps: int = 2):
        """Smoothed values at every x.  `nsteps` is an extension: the reference
        hard-codes 2 robustness iterations (:866)."""
        errmgr = err if err is not None else errors()
        if not self.m_init:                                   # :850-855
            self._report(errmgr, "ls_smooth", CF_NO_DATA_DEFINED_ERROR)
            return None
        n = self.m_x.size
        ys = np.empty(n)                                      # :859
        st = _lib.cf_stats()
        rc = self._lib.cf_lowess_run(self._h, float(f), int(nsteps), C.byref(st))
        if rc == 0:
            if n < 2:                                         # :651-654: rw, res untouched
                rc = self._lib.cf_lowess_fetch(self._h, _dp(ys), None, None)
            else:
                rc = self._lib.cf_lowess_fetch(self._h, _dp(ys),
                                               _dp(self.m_weights) if nsteps > 0 else None,
                                               _dp(self.m_residuals))
        if rc != 0:
            self._report(errmgr, "ls_smooth", rc)
            return None
        self.last_stats = st
        return ys

    # -- device-resident run (benchmark path; no host copies) -------------------------
    def run_resident(self, f: float, nsteps: int = 2, fresh_windows: bool = True) -> "_lib.cf_stats":
        """One complete smooth() on the resident data.  fresh_windows: recompute the
        neighbourhoods like every call of the reference's lowess does."""
        if not self.m_init:
            raise RuntimeError("No data has been defined.")
        if fresh_windows:
            self._lib.cf_lowess_reset_windows(self._h)
        st = _lib.cf_stats()
        rc = self._lib.cf_lowess_run(self._h, float(f), int(nsteps), C.byref(st))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_run failed ({rc}): {_lib.last_error()}")
        self.last_stats = st
        return st

    def windows(self, f: float):
        """1-based neighbourhood indices for span fraction f (index-parity tests)."""
        if not self.m_init:
            raise RuntimeError("No data has been defined.")
        n = self.m_x.size
        names = ("nleft", "nright", "lo9", "hi9", "lo1", "hi1", "leader")
        arrs = {k: np.zeros(n, np.int32) for k in names}
        rc = self._lib.cf_lowess_windows(self._h, float(f), *[_ip(arrs[k]) for k in names])
        if rc != 0:
            raise RuntimeError(f"cf_lowess_windows failed ({rc}): {_lib.last_error()}")
        return arrs

    def run_from_rw(self, f: float, rw, passes: int = 1):
        """Teacher forcing: `passes` fit passes starting from robustness weights rw
        (None = unweighted first pass).  Returns (ys, rw, res)."""
        if not self.m_init:
            raise RuntimeError("No data has been defined.")
        n = self.m_x.size
        rwc = None if rw is None else np.ascontiguousarray(rw, dtype=np.float64)
        rc = self._lib.cf_lowess_run_from_rw(self._h, float(f), _dp(rwc), int(passes))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_run_from_rw failed ({rc}): {_lib.last_error()}")
        ys, rwo, res = np.empty(n), np.empty(n), np.empty(n)
        rc = self._lib.cf_lowess_fetch(self._h, _dp(ys), _dp(rwo), _dp(res))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fetch failed ({rc}): {_lib.last_error()}")
        return ys, rwo, res

    def fit_sample(self, f: float, use_rw: bool, first_tile: int, tile_stride: int, ntiles: int,
                   tile: int = 512):
        """One pass of the DIRECT kernel over the strided tile sample (cf_lowess_fit_sample).
        Returns (ms of the launch, 0-based indices of the sampled points, ys at them)."""
        if not self.m_init:
            raise RuntimeError("No data has been defined.")
        n = self.m_x.size
        ms = C.c_float(0.0)
        rc = self._lib.cf_lowess_fit_sample(self._h, float(f), 1 if use_rw else 0, int(first_tile),
                                            int(tile_stride), int(ntiles), C.byref(ms))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fit_sample failed ({rc}): {_lib.last_error()}")
        b, e = C.c_int32(), C.c_int32()
        self._lib.cf_lowess_shard(self._h, C.byref(b), C.byref(e), None)   # tiles count from the shard's start
        tiles_all = -(-(e.value - b.value) // tile)
        t = np.arange(first_tile, tiles_all, tile_stride, dtype=np.int64)[:ntiles]
        idx = (b.value + t[:, None] * tile + np.arange(tile)[None, :]).ravel()
        idx = idx[idx < e.value]
        ys = np.empty(n)
        rc = self._lib.cf_lowess_fetch(self._h, _dp(ys), None, None)
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fetch failed ({rc}): {_lib.last_error()}")
        return float(ms.value), idx, ys[idx]

    # -- accessors (:876-939) ------------------------------------------------------------
    def get_count(self) -> int:
        return int(self.m_x.size) if self.m_init else 0       # :876-884

    def get_x(self, ind: int) -> float:
        return float(self.m_x[ind - 1]) if self.m_init else 0.0   # :893-902 (1-based)

    def get_y(self, ind: int) -> float:
        return float(self.m_y[ind - 1]) if self.m_init else 0.0   # :911-920

    def get_residuals(self, x: np.ndarray) -> None:
        n = min(x.size, self.get_count())                     # :937
        x[:n] = self.m_residuals[:n]                          # :938


class batched_lowess:
    """nseries independent lowess_smoothing objects of equal length held on one GPU
    and smoothed together (series are a grid dimension of every kernel).  Not part
    of the reference's API: it is BASELINE config 3's device-resident form."""

    def __init__(self, device: int = -1, stream: int = 0):
        self._lib = _lib.load()
        self._h = C.c_void_p()
        rc = self._lib.cf_lowess_create(C.byref(self._h), device, C.c_void_p(stream) if stream else None)
        if rc != 0:
            self._h = C.c_void_p()
            raise RuntimeError(f"cf_lowess_create failed ({rc}): {_lib.last_error()}")
        self.shape = (0, 0)

    def set_method(self, method) -> None:
        code = _METHODS[method] if isinstance(method, str) else int(method)
        rc = self._lib.cf_lowess_set_method(self._h, code)
        if rc != 0:
            raise RuntimeError(f"cf_lowess_set_method failed ({rc}): {_lib.last_error()}")

    def initialize(self, x, y) -> None:
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.ascontiguousarray(y, dtype=np.float64)
        assert x.ndim == 2 and x.shape == y.shape
        rc = self._lib.cf_lowess_init_batched(self._h, _dp(x), _dp(y), x.shape[0], x.shape[1])
        if rc != 0:
            raise RuntimeError(f"cf_lowess_init_batched failed ({rc}): {_lib.last_error()}")
        self.shape = x.shape

    def run_resident(self, f: float, nsteps: int = 2, fresh_windows: bool = True) -> "_lib.cf_stats":
        if fresh_windows:
            self._lib.cf_lowess_reset_windows(self._h)
        st = _lib.cf_stats()
        rc = self._lib.cf_lowess_run(self._h, float(f), int(nsteps), C.byref(st))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_run failed ({rc}): {_lib.last_error()}")
        return st

    def fetch(self):
        ys, rw, res = (np.empty(self.shape) for _ in range(3))
        rc = self._lib.cf_lowess_fetch(self._h, _dp(ys), _dp(rw), _dp(res))
        if rc != 0:
            raise RuntimeError(f"cf_lowess_fetch failed ({rc}): {_lib.last_error()}")
        return ys, rw, res

    def close(self) -> None:
        if self._h:
            self._lib.cf_lowess_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- stateless wrappers over the host-buffer ABI (what the Fortran shim binds) --------
def cf_sort_xy(x, y):
    lib = _lib.load()
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = x.size
    xs, ys, perm = np.empty(n), np.empty(n), np.zeros(n, np.int32)
    rc = lib.cf_sort_xy(_dp(x), _dp(y), n, _dp(xs), _dp(ys), _ip(perm))
    if rc != 0:
        raise RuntimeError(f"cf_sort_xy failed ({rc}): {_lib.last_error()}")
    return xs, ys, perm


def cf_lowess_smooth(x_sorted, y, f, nsteps=2, out=None):
    """lowess(x, y, f, nsteps, 0, ys, rw, res) through the host-buffer C ABI.
    `out` may hold preallocated (ys, rw, res) arrays (e.g. pinned memory)."""
    lib = _lib.load()
    x = np.ascontiguousarray(x_sorted, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = x.size
    if out is None:
        out = (np.empty(n), np.zeros(n), np.zeros(n))
    ys, rw, res = out
    rc = lib.cf_lowess_smooth(_dp(x), _dp(y), n, float(f), int(nsteps), _dp(ys), _dp(rw), _dp(res))
    if rc != 0:
        raise RuntimeError(f"cf_lowess_smooth failed ({rc}): {_lib.last_error()}")
    return ys, rw, res


def cf_lowess_smooth_multi(x_sorted, y, f, nsteps=2, devices=None, out=None):
    """cf_lowess_smooth over several GPUs of the box from this one process (csrc/multi.cu):
    the series is range-sharded over `devices` (None = every visible device), each device
    receives only its block + halo and writes its own block of the result."""
    lib = _lib.load()
    x = np.ascontiguousarray(x_sorted, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = x.size
    if out is None:
        out = (np.empty(n), np.zeros(n), np.zeros(n))
    ys, rw, res = out
    if devices is None:
        devs, ndev = None, 0
    else:
        devs = (C.c_int32 * len(devices))(*[int(d) for d in devices])
        ndev = len(devices)
    rc = lib.cf_lowess_smooth_multi(_dp(x), _dp(y), n, float(f), int(nsteps), devs, ndev, _dp(ys), _dp(rw), _dp(res))
    if rc != 0:
        raise RuntimeError(f"cf_lowess_smooth_multi failed ({rc}): {_lib.last_error()}")
    return ys, rw, res


def cf_lowess_smooth_delta(x_sorted, y, f, nsteps=2, delta=0.0):
    """lowess(x, y, f, nsteps, delta, ys, rw, res) (:625-715) through the host-buffer C ABI."""
    lib = _lib.load()
    x = np.ascontiguousarray(x_sorted, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = x.size
    ys, rw, res = np.empty(n), np.zeros(n), np.zeros(n)
    rc = lib.cf_lowess_smooth_delta(_dp(x), _dp(y), n, float(f), int(nsteps), float(delta),
                                    _dp(ys), _dp(rw), _dp(res))
    if rc != 0:
        raise RuntimeError(f"cf_lowess_smooth_delta failed ({rc}): {_lib.last_error()}")
    return ys, rw, res


def cf_lowess_smooth_batched(x_sorted, y, f, nsteps=2):
    lib = _lib.load()
    x = np.ascontiguousarray(x_sorted, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    assert x.ndim == 2 and x.shape == y.shape
    s, n = x.shape
    ys, rw, res = np.empty((s, n)), np.zeros((s, n)), np.zeros((s, n))
    rc = lib.cf_lowess_smooth_batched(_dp(x), _dp(y), s, n, float(f), int(nsteps),
                                      _dp(ys), _dp(rw), _dp(res))
    if rc != 0:
        raise RuntimeError(f"cf_lowess_smooth_batched failed ({rc}): {_lib.last_error()}")
    return ys, rw, res


def fp64_peak(device: int = -1, ms_target: float = 200.0):
    lib = _lib.load()
    tf, mhz = C.c_double(), C.c_double()
    rc = lib.cf_fp64_peak(device, ms_target, C.byref(tf), C.byref(mhz))
    if rc != 0:
        raise RuntimeError(f"cf_fp64_peak failed ({rc}): {_lib.last_error()}")
    return tf.value, mhz.value
