"""Host-side mirror of the one curvefit_statistics routine that shares machinery with
the LOWESS path: `median` (src/curvefit_statistics.f90:258-295).  The numerical work
runs in libcurvefit_cuda (exact radix select of the middle order statistics; radix sort
for the srt = .true. side effect)."""
from __future__ import annotations

import numpy as np

from . import _lib


def median(x: np.ndarray, srt: bool = True) -> float:
    """median(x [, srt]).  Like the reference, srt = True (the default) leaves x sorted
    into ascending order on return (:286); srt = False asserts that x is already
    ascending and reads the middle element(s) as they lie (:289-293)."""
    if not (isinstance(x, np.ndarray) and x.dtype == np.float64 and x.flags.c_contiguous and x.ndim == 1):
        raise TypeError("x must be a contiguous 1-d float64 array (it may be sorted in place)")
    lib = _lib.load()
    n = x.size
    if n == 0:
        return 0.0                                            # :280-283
    if not srt:
        nmid = n // 2                                         # :273-275, 1-based nmid, nmid+1
        return float(x[nmid]) if n % 2 else float(0.5 * (x[nmid - 1] + x[nmid]))
    z = np.zeros(1)
    rc = lib.cf_median(x.ctypes.data_as(_lib.c_dp), n, z.ctypes.data_as(_lib.c_dp))
    if rc == 0:
        out = np.empty(n)
        rc = lib.cf_sort(x.ctypes.data_as(_lib.c_dp), n, out.ctypes.data_as(_lib.c_dp))
        if rc == 0:
            x[:] = out
    if rc != 0:
        raise RuntimeError(f"median failed ({rc}): {_lib.last_error()}")
    return float(z[0])
