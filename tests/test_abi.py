"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every
symbol include/curvefit_cuda.h declares, and fails LOUDLY without a GPU (no
CPU fallback).  No device compute here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from curvefit_b200 import _lib, core
from curvefit_b200.errors import errors
from curvefit_b200.regression import lowess_smoothing

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "curvefit_cuda.h")).read()
    return sorted(set(re.findall(r"CF_API\s+[\w\s\*]+?\b(cf_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = _lib.load()
    names = _declared_symbols()
    assert len(names) >= 25
    for name in names:
        assert hasattr(lib, name), name
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    assert lib.cf_version() >= 100


def test_span_matches_reference_truncation():
    lib = _lib.load()
    assert lib.cf_lowess_span(16384, 0.3) == 4915        # F3: trunc, not ceil
    assert lib.cf_lowess_span(100000000, 0.05) == 5000000
    assert lib.cf_lowess_span(100, 0.2) == 20
    assert lib.cf_lowess_span(100, 0.0) == 2
    assert lib.cf_lowess_span(100, 7.0) == 100


def test_error_codes_are_the_references():
    # /root/reference src/curvefit_core.f90:34-40
    assert (core.CF_ARRAY_SIZE_ERROR, core.CF_OUT_OF_MEMORY_ERROR,
            core.CF_NO_DATA_DEFINED_ERROR, core.CF_INVALID_INPUT_ERROR) == (202, 105, 300, 201)
    hdr = open(os.path.join(ROOT, "include", "curvefit_cuda.h")).read()
    for name, val in (("CF_ARRAY_SIZE_ERROR", 202), ("CF_OUT_OF_MEMORY_ERROR", 105),
                      ("CF_NO_DATA_DEFINED_ERROR", 300), ("CF_INVALID_INPUT_ERROR", 201),
                      ("CF_DEVICE_ERROR", 303)):
        assert re.search(rf"#define\s+{name}\s+{val}\b", hdr), name


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_without_a_device():
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.cf_lowess_create(C.byref(h), -1, None) == core.CF_DEVICE_ERROR
    assert "no CPU fallback" in _lib.last_error() or "CUDA" in _lib.last_error()
    x = np.linspace(0, 1, 10)
    ys = np.zeros(10)
    rc = lib.cf_lowess_smooth(x.ctypes.data_as(_lib.c_dp), x.ctypes.data_as(_lib.c_dp), 10, 0.5, 2,
                              ys.ctypes.data_as(_lib.c_dp), ys.ctypes.data_as(_lib.c_dp),
                              ys.ctypes.data_as(_lib.c_dp))
    assert rc == core.CF_DEVICE_ERROR
    fit = lowess_smoothing()
    with pytest.raises(core.CurvefitError) as ei:
        fit.initialize(x, x)
    assert ei.value.flag == core.CF_DEVICE_ERROR and ei.value.fcn == "ls_init"


def test_host_side_error_behaviour_mirrors_reference():
    """ls_init size check (:769-773) and ls_smooth's m_init guard (:850-855) with
    a caller-supplied error object that does not terminate."""
    fit = lowess_smoothing()
    err = errors()
    err.set_exit_on_error(False)
    err.set_suppress_printing(True)
    fit.initialize(np.zeros(5), np.zeros(4), err=err)
    assert err.has_error_occurred() and err.get_error_flag() == 202
    assert err.get_error_fcn_name() == "ls_init"
    assert err.get_error_message() == "Input array sizes must match."
    assert fit.get_count() == 0 and fit.get_x(1) == 0.0 and fit.get_y(3) == 0.0
    err.reset_error_status()
    assert fit.smooth(0.5, err=err) is None
    assert err.get_error_flag() == 300 and err.get_error_fcn_name() == "ls_smooth"
    assert err.get_error_message() == "No data has been defined."
    # default error object "terminates": here it raises
    with pytest.raises(core.CurvefitError) as ei:
        lowess_smoothing().smooth(0.5)
    assert ei.value.flag == 300


def test_moving_average_and_median_host_checks_need_no_device():
    """moving_average_1's input check (:452-457) and median's n = 0 / srt = .false. paths
    (src/curvefit_statistics.f90:280-293) are host logic; the device entry points fail loudly
    without a GPU."""
    import curvefit_b200 as cf
    x = np.arange(6.0)
    for npts in (1, 7):
        err = errors()
        err.set_exit_on_error(False)
        err.set_suppress_printing(True)
        cf.moving_average(x, npts, err=err)
        assert err.get_error_flag() == core.CF_INVALID_INPUT_ERROR
        assert err.get_error_fcn_name() == "moving_average_1"
    assert np.array_equal(x, np.arange(6.0))
    with pytest.raises(core.CurvefitError):
        cf.moving_average(x, 1)                      # default error object "terminates"
    assert cf.median(np.zeros(0)) == 0.0
    assert cf.median(np.array([1.0, 2.0, 4.0, 8.0]), srt=False) == 3.0
    assert cf.median(np.array([1.0, 2.0, 4.0]), srt=False) == 2.0
    lib = _lib.load()
    assert lib.cf_moving_average(x.ctypes.data_as(_lib.c_dp), 6, 1) == core.CF_INVALID_INPUT_ERROR
    if not _have_gpu():
        assert lib.cf_moving_average(x.ctypes.data_as(_lib.c_dp), 6, 3) == core.CF_DEVICE_ERROR
        z = C.c_double()
        assert lib.cf_median(x.ctypes.data_as(_lib.c_dp), 6, C.byref(z)) == core.CF_DEVICE_ERROR
        out = np.zeros(6)
        assert lib.cf_sort(x.ctypes.data_as(_lib.c_dp), 6, out.ctypes.data_as(_lib.c_dp)) == core.CF_DEVICE_ERROR
        assert lib.cf_lowess_smooth_delta(x.ctypes.data_as(_lib.c_dp), x.ctypes.data_as(_lib.c_dp), 6, 0.5, 2, 0.1,
                                          out.ctypes.data_as(_lib.c_dp), out.ctypes.data_as(_lib.c_dp),
                                          out.ctypes.data_as(_lib.c_dp)) == core.CF_DEVICE_ERROR


# ---- the boundary from a compiled language (the reference is compiled code) ---------------
def _cc(args, **kw):
    import subprocess
    return subprocess.run(args, cwd=ROOT, capture_output=True, text=True, **kw)


@pytest.mark.parametrize("compiler,std", [("gcc", "-std=c99"), ("g++", "-std=c++17")])
def test_header_is_plain_c_and_cxx(compiler, std, tmp_path):
    """include/curvefit_cuda.h: no torch / C++ types, compiles pedantically as C99 and as C++."""
    src = tmp_path / ("t.c" if compiler == "gcc" else "t.cpp")
    src.write_text('#include "curvefit_cuda.h"\nint main(void) { return cf_version() > 0 ? 0 : 1; }\n')
    r = _cc([compiler, std, "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only",
             "-I", os.path.join(ROOT, "include"), str(src)])
    assert r.returncode == 0, r.stderr


def _build_example(tmp_path):
    exe = str(tmp_path / "lowess_c_abi")
    libdir = os.path.join(ROOT, "curvefit_b200")
    r = _cc(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-Iinclude",
             "examples/lowess_c_abi.c", "-L" + libdir, "-lcurvefit_cuda", "-Wl,-rpath," + libdir,
             "-lm", "-o", exe])
    assert r.returncode == 0, r.stderr
    return exe


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_c_example_links_and_fails_loudly_without_a_device(tmp_path):
    _lib.load()                                   # ImportError if the library has not been built
    r = _cc([_build_example(tmp_path)], timeout=120)
    assert r.returncode == 3, (r.returncode, r.stdout, r.stderr)
    assert "code 303" in r.stderr and "no CPU fallback" in r.stderr


@pytest.mark.gpu
def test_c_example_runs_on_the_device(tmp_path):
    _lib.load()
    r = _cc([_build_example(tmp_path)], timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "ys =" in r.stdout


def test_cmake_build_produces_the_same_abi(tmp_path):
    """north_star names the repo's CMake as the build: configure + build into a scratch
    directory (sm_100a only) and check that the library exports every symbol the header
    declares.  (build() uses the Makefile; both compile the same sources with the same flags.)"""
    import shutil
    import subprocess
    cmake = shutil.which("cmake")
    if cmake is None or shutil.which("nvcc") is None:
        pytest.skip("cmake / nvcc not on PATH")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    bdir, odir = str(tmp_path / "b"), str(tmp_path / "out")
    subprocess.run([cmake, "-S", root, "-B", bdir, "-DCMAKE_CUDA_HOST_COMPILER=/usr/bin/g++",
                    "-DCMAKE_CXX_COMPILER=/usr/bin/g++", f"-DCURVEFIT_OUTPUT_DIR={odir}"],
                   check=True, env=env, capture_output=True, timeout=300)
    subprocess.run([cmake, "--build", bdir, "-j8"], check=True, env=env, capture_output=True, timeout=900)
    so = os.path.join(odir, "libcurvefit_cuda.so")
    assert os.path.exists(so)
    syms = subprocess.run(["nm", "-D", "--defined-only", so], check=True, capture_output=True, text=True).stdout
    have = {ln.split()[-1] for ln in syms.splitlines() if ln.strip()}
    from curvefit_b200 import _lib
    missing = [name for name in _lib.SIGNATURES if name not in have]
    assert not missing, missing
    sass = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    assert "sm_100a" in sass and "sm_90" not in sass


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal", "wide"])
@pytest.mark.parametrize("n,f", [(3000, 0.1), (5000, 0.6), (777, 0.02)])
def test_block_span_covers_every_window_of_the_block(kind, n, f):
    """cf_lowess_block_span (host-only: what a rank of a range-sharded series uploads) against the
    oracle's recorded windows: for every block of output points the span must contain each point's
    nleft .. nrt (the sweep stops ON the first excluded point, :553-554), its tie-group leader and
    the extra element the window search tests, and it must not exceed the (ns-1)-point halo."""
    from oracle import oracle
    from tests import cases
    oracle.build()
    lib = _lib.load()
    x = cases.make_x(kind, n, 11)
    y = cases.make_y(x, 11)
    ns = max(min(int(f * float(n)), n), 2)
    r = oracle.lowess(x, y, f, 0, want_windows=True)
    lead = np.searchsorted(x, x, side="left")                       # 0-based leader of every point
    nl = r["nleft"][lead] - 1                                        # followers take their leader's window
    nrt = r["nrt"][lead] - 1
    nr = r["nright"][lead] - 1
    p = lambda a: a.ctypes.data_as(_lib.c_dp)
    for parts in (1, 2, 3, 7):
        edges = np.linspace(0, n, parts + 1).astype(int)
        for b, e in zip(edges[:-1], edges[1:]):
            if e <= b:
                continue
            first, last = C.c_int32(), C.c_int32()
            rc = lib.cf_lowess_block_span(p(x), n, f, int(b) + 1, int(e), C.byref(first), C.byref(last))
            assert rc == 0
            lo, hi = first.value - 1, last.value - 1
            assert lo <= min(nl[b:e].min(), lead[b:e].min()), (kind, b, e)
            assert hi >= min(max(nrt[b:e].max(), nr[b:e].max() + 1), n - 1), (kind, b, e)
            if kind not in ("ties", "equal"):                        # ties widen it legitimately
                assert lo >= b - ns and hi <= e - 1 + ns, (kind, b, e)
    assert lib.cf_lowess_block_span(p(x), n, f, 0, 5, C.byref(first), C.byref(last)) == 201
