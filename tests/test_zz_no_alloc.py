"""cf_median_device_noalloc: the median for callers that may not allocate device memory.

Some users of the library cannot let it call cudaMalloc / cudaMallocAsync: another allocator in
their process already owns the whole GPU, their device has no stream-ordered memory pools, or they
are past their set-up phase and may not take the driver's allocation locks any more.  For them the
select's scratch is a fixed array inside the library (at most CF_MEDIAN_NOALLOC_MAX_N = 1 048 576
elements).  The entry point has no counterpart in the reference; its contract is "the same answer
as cf_median_device", which is what this file checks -- bit for bit, the tolerance of the other
median tests -- on two calls in a row, because every call reuses the one array and must not be
disturbed by what the previous call left there.

The file is named so that pytest collects it last: with `-x` a failure here hides nothing.
It must pass alone in a fresh process (the first call does the per-device set-up) and after the
rest of the suite (the set-up may already have happened; other kernels have run).
"""
import ctypes as C
import re
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "curvefit_cuda.h")


def _max_n():
    text = open(HEADER).read()
    return int(re.search(r"#define\s+CF_MEDIAN_NOALLOC_MAX_N\s+(\d+)", text).group(1))


@pytest.fixture(scope="module")
def lib():
    from curvefit_b200 import _lib
    return _lib.load()


def _both(lib, torch, x):
    """(existing entry point, no-alloc entry point) on the same device array, default stream."""
    n = x.numel()
    out = torch.full((2,), float("nan"), dtype=torch.float64, device="cuda")
    ptr = out.data_ptr()
    assert lib.cf_median_device(C.c_void_p(x.data_ptr()), n, C.c_void_p(ptr), None) == 0
    assert lib.cf_median_device_noalloc(C.c_void_p(x.data_ptr()), n, C.c_void_p(ptr + 8), None) == 0
    torch.cuda.synchronize()
    return float(out[0].item()), float(out[1].item())


def test_two_calls_in_a_row_match_the_allocating_entry_point(lib):
    import torch
    rng = np.random.default_rng(2024)
    # the second series is shorter and has a different scale, so every counter and every candidate
    # slot the first call left behind would show if it were reused without clearing
    first = rng.normal(size=300_001) * np.exp(rng.normal(size=300_001) * 4.0)
    second = 1.0 + rng.integers(0, 1000, 4096) * np.finfo(np.float64).eps
    for host in (first, second):
        x = torch.from_numpy(host).cuda()
        want, got = _both(lib, torch, x)
        s = np.sort(host)
        n = host.size
        ref = s[n // 2] if n % 2 else 0.5 * (s[n // 2 - 1] + s[n // 2])      # median_dbl :280-292
        assert got == want == ref, (n, got, want, ref)


def test_same_input_twice_and_the_largest_size(lib):
    import torch
    n = _max_n()
    rng = np.random.default_rng(7)
    host = rng.normal(size=n)
    host[rng.integers(0, n, n // 8)] = 0.25                                   # ties
    x = torch.from_numpy(host).cuda()
    a = _both(lib, torch, x)
    b = _both(lib, torch, x)
    assert a[0] == a[1] == b[0] == b[1]
    for m in (1, 2, 3):                                                        # smallest sizes after the largest
        want, got = _both(lib, torch, x[:m].contiguous())
        assert got == want


def test_sizes_beyond_the_built_in_scratch_are_refused(lib):
    import torch
    from curvefit_b200 import CF_ARRAY_SIZE_ERROR
    n = _max_n() + 1
    x = torch.zeros(n, dtype=torch.float64, device="cuda")
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    rc = lib.cf_median_device_noalloc(C.c_void_p(x.data_ptr()), n, C.c_void_p(out.data_ptr()), None)
    assert rc == CF_ARRAY_SIZE_ERROR
    buf = C.create_string_buffer(256)
    lib.cf_last_error(buf, 256)
    assert b"built-in scratch" in buf.value and str(_max_n()).encode() in buf.value
    torch.cuda.synchronize()
    assert float(out.item()) == -1.0                                           # nothing was launched
    assert lib.cf_median_device_noalloc(None, 5, C.c_void_p(out.data_ptr()), None) == CF_ARRAY_SIZE_ERROR
