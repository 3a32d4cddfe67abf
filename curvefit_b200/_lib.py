"""ctypes loader for libcurvefit_cuda.so (the C ABI in include/curvefit_cuda.h).

The library is built in-tree (Makefile / CMakeLists.txt / __graft_entry__.build)
as curvefit_b200/libcurvefit_cuda.so.  Loading fails loudly when it is missing:
there is no Python or CPU implementation to fall back to.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libcurvefit_cuda.so")

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)


class cf_stats(C.Structure):
    _fields_ = [
        ("ms_total", C.c_float),
        ("ms_windows", C.c_float),
        ("ms_fit", C.c_float),
        ("ms_robust", C.c_float),
        ("fit_launches", C.c_int32),
        ("launches", C.c_int32),
        ("ns", C.c_int32),
        ("passes", C.c_int32),
        ("cmad_last", C.c_double),
        ("ms_tree_build", C.c_float),
        ("reserved", C.c_float),
    ]


# name -> (restype, argtypes); every symbol declared in include/curvefit_cuda.h
SIGNATURES = {
    "cf_version": (C.c_int32, []),
    "cf_last_error": (C.c_int32, [C.c_char_p, C.c_int32]),
    "cf_lowess_span": (C.c_int32, [C.c_int32, C.c_double]),
    "cf_sort_xy": (C.c_int32, [c_dp, c_dp, C.c_int32, c_dp, c_dp, c_ip]),
    "cf_lowess_smooth": (C.c_int32, [c_dp, c_dp, C.c_int32, C.c_double, C.c_int32, c_dp, c_dp, c_dp]),
    "cf_lowess_smooth_ex": (C.c_int32, [c_dp, c_dp, C.c_int32, C.c_double, C.c_int32, C.c_int32, c_dp, c_dp, c_dp]),
    "cf_lowess_smooth_delta": (C.c_int32, [c_dp, c_dp, C.c_int32, C.c_double, C.c_int32, C.c_double,
                                           c_dp, c_dp, c_dp]),
    "cf_lowess_set_delta": (C.c_int32, [C.c_void_p, C.c_double]),
    "cf_lowess_delta": (C.c_double, [C.c_void_p]),
    "cf_lowess_fitted": (C.c_int32, [C.c_void_p, c_ip, c_ip, C.c_int32]),
    "cf_lowess_smooth_batched": (C.c_int32, [c_dp, c_dp, C.c_int32, C.c_int32, C.c_double, C.c_int32,
                                             c_dp, c_dp, c_dp]),
    "cf_lowess_smooth_multi": (C.c_int32, [c_dp, c_dp, C.c_int32, C.c_double, C.c_int32, c_ip, C.c_int32,
                                           c_dp, c_dp, c_dp]),
    "cf_lowess_multi_release": (C.c_int32, []),
    "cf_lowess_multi_active": (C.c_int32, []),
    "cf_moving_average": (C.c_int32, [c_dp, C.c_int32, C.c_int32]),
    "cf_moving_average_device": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "cf_median": (C.c_int32, [c_dp, C.c_int32, c_dp]),
    "cf_median_device": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "cf_median_device_noalloc": (C.c_int32, [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "cf_sort": (C.c_int32, [c_dp, C.c_int32, c_dp]),
    "cf_lowess_create": (C.c_int32, [C.POINTER(C.c_void_p), C.c_int32, C.c_void_p]),
    "cf_lowess_destroy": (None, [C.c_void_p]),
    "cf_lowess_init": (C.c_int32, [C.c_void_p, c_dp, c_dp, C.c_int32, C.c_int32]),
    "cf_lowess_init_device": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "cf_lowess_shard_span": (C.c_int32, [C.c_void_p, c_dp, C.c_int32, C.c_double, c_ip, c_ip]),
    "cf_lowess_block_span": (C.c_int32, [c_dp, C.c_int32, C.c_double, C.c_int32, C.c_int32, c_ip, c_ip]),
    "cf_lowess_init_span": (C.c_int32, [C.c_void_p, c_dp, c_dp, C.c_int32, C.c_int32, C.c_int32]),
    "cf_lowess_fetch_block": (C.c_int32, [C.c_void_p, c_dp, c_dp, c_dp]),
    "cf_lowess_set_early_weights": (C.c_int32, [C.c_void_p, c_dp]),
    "cf_lowess_init_batched": (C.c_int32, [C.c_void_p, c_dp, c_dp, C.c_int32, C.c_int32]),
    "cf_lowess_set_shard": (C.c_int32, [C.c_void_p, C.c_int32, C.c_int32]),
    "cf_lowess_shard": (C.c_int32, [C.c_void_p, c_ip, c_ip, c_ip]),
    "cf_lowess_run": (C.c_int32, [C.c_void_p, C.c_double, C.c_int32, C.POINTER(cf_stats)]),
    "cf_lowess_run_from_rw": (C.c_int32, [C.c_void_p, C.c_double, c_dp, C.c_int32]),
    "cf_lowess_fit_sample": (C.c_int32, [C.c_void_p, C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                         C.POINTER(C.c_float)]),
    "cf_lowess_fetch": (C.c_int32, [C.c_void_p, c_dp, c_dp, c_dp]),
    "cf_lowess_fetch_xy": (C.c_int32, [C.c_void_p, c_dp, c_dp]),
    "cf_lowess_count": (C.c_int32, [C.c_void_p]),
    "cf_lowess_launch_count": (C.c_int64, [C.c_void_p]),
    "cf_lowess_set_method": (C.c_int32, [C.c_void_p, C.c_int32]),
    "cf_lowess_method": (C.c_int32, [C.c_void_p]),
    "cf_set_default_method": (C.c_int32, [C.c_int32]),
    "cf_default_method": (C.c_int32, []),
    "cf_lowess_reset_windows": (C.c_int32, [C.c_void_p]),
    "cf_lowess_windows": (C.c_int32, [C.c_void_p, C.c_double, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip, c_ip]),
    "cf_lowess_stage_begin": (C.c_int32, [C.c_void_p, C.c_double]),
    "cf_lowess_stage_fit": (C.c_int32, [C.c_void_p, C.c_int32]),
    "cf_lowess_stage_hist": (C.c_int32, [C.c_void_p, C.c_int32]),
    "cf_lowess_stage_pick": (C.c_int32, [C.c_void_p, C.c_int32]),
    "cf_lowess_stage_weights": (C.c_int32, [C.c_void_p]),
    "cf_lowess_select_rounds": (C.c_int32, []),
    "cf_lowess_peer_handle_bytes": (C.c_int32, []),
    "cf_lowess_peer_export": (C.c_int32, [C.c_void_p, C.c_void_p]),
    "cf_lowess_peer_attach": (C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32]),
    "cf_lowess_peer_attach_local": (C.c_int32, [C.POINTER(C.c_void_p), C.c_int32]),
    "cf_lowess_peer_set_timeout": (C.c_int32, [C.c_void_p, C.c_double]),
    "cf_lowess_peer_detach": (C.c_int32, [C.c_void_p]),
    "cf_lowess_peer_attached": (C.c_int32, [C.c_void_p]),
    "cf_lowess_peer_error": (C.c_int32, [C.c_void_p, c_ip]),
    "cf_lowess_stage_gather": (C.c_int32, [C.c_void_p, C.c_int32]),
    "cf_lowess_sync": (C.c_int32, [C.c_void_p]),
    "cf_lowess_buffer": (C.c_void_p, [C.c_void_p, C.c_int32]),
    "cf_lowess_hist_len": (C.c_int32, []),
    "cf_lowess_hist0_len": (C.c_int32, []),
    "cf_fp64_peak": (C.c_int32, [C.c_int32, C.c_double, c_dp, c_dp]),
}

CF_GATHER_YS, CF_GATHER_RES, CF_GATHER_RW = 1, 2, 4
CF_METHOD_DIRECT, CF_METHOD_TREE, CF_METHOD_AUTO = range(3)
CF_BUF_X, CF_BUF_Y, CF_BUF_YS, CF_BUF_RW, CF_BUF_RES, CF_BUF_HIST, CF_BUF_CMAD, CF_BUF_HIST0 = range(8)

_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; raises ImportError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make` (or "
                "`python -c 'import __graft_entry__ as g; g.build()'`). "
                "curvefit_b200 has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the ABI is incomplete
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error() -> str:
    buf = C.create_string_buffer(512)
    load().cf_last_error(buf, 512)
    return buf.value.decode("utf-8", "replace")
