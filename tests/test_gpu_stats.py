"""GPU parity of the LOWESS path's neighbours against the oracle:
moving_average bit-exact (same recurrence, IEEE operations), median / sort bit-exact
(order statistics and sorted values do not depend on the algorithm)."""
import numpy as np
import pytest

from curvefit_b200 import datagen

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cf():
    import curvefit_b200
    return curvefit_b200


@pytest.mark.parametrize("n,npts", [(2, 2), (10, 3), (100, 2), (100, 100), (1000, 17), (5000, 257),
                                    (4096, 256), (4097, 258), (20000, 1000), (300, 299)])
def test_moving_average_bit_exact(cf, oracle, n, npts):
    rng = np.random.default_rng(n + npts)
    x = datagen.noisy_y(datagen.grid_x(n), npts) * rng.choice([1.0, 1e6, 1e-6])
    ref, rc = oracle.moving_average(x, npts)
    assert rc == 0
    got = x.copy()
    cf.moving_average(got, npts)
    assert np.array_equal(got, ref)


def test_moving_average_example_shape(cf, oracle):
    """examples/curvefit_moving_avg_example.f90: 1000-point noisy signal, 10-point window."""
    n = 1000
    x = np.linspace(0.0, 1.0, n)
    y = np.sin(10.0 * x) + (datagen.uniform01(3, n) - 0.5) * 0.5
    ref, _ = oracle.moving_average(y, 10)
    got = y.copy()
    cf.moving_average(got, 10)
    assert np.array_equal(got, ref)
    # zeros are averaged in at the start (:467-473)
    assert abs(got[0]) <= abs(y[0]) * 0.11


def test_moving_average_errors(cf):
    from curvefit_b200 import CF_INVALID_INPUT_ERROR, errors
    x = np.arange(10.0)
    keep = x.copy()
    for npts in (1, 0, -3, 11):
        e = errors()
        e.set_exit_on_error(False)
        e.set_suppress_printing(True)
        cf.moving_average(x, npts, err=e)
        assert e.has_error_occurred() and e.get_error_flag() == CF_INVALID_INPUT_ERROR
        assert np.array_equal(x, keep)


def test_median_known_answer(cf):
    """tests/curvefit_test_statistics.f90:286-309."""
    x = np.array([2.0, 3.0, 1.0, -4.0, -7.0, 1.0, 3.0])
    assert cf.median(x) == 1.0
    assert np.array_equal(x, np.array([-7.0, -4.0, 1.0, 1.0, 2.0, 3.0, 3.0]))   # srt = .true. sorts x


@pytest.mark.parametrize("n", [1, 2, 3, 10, 1001, 4096, 50001, 1 << 20])
def test_median_and_sort_bit_exact(cf, oracle, n):
    rng = np.random.default_rng(n)
    x = rng.normal(size=n) * np.exp(rng.normal(size=n) * 5.0)
    if n > 10:
        x[rng.integers(0, n, n // 10)] = 0.0
        x[rng.integers(0, n, n // 10)] = x[0]
    z_ref, sorted_ref = oracle.median(x, True)
    a = x.copy()
    z = cf.median(a)
    assert z == z_ref
    assert np.array_equal(a, sorted_ref)
    assert cf.median(a, srt=False) == z_ref


def test_median_empty(cf):
    assert cf.median(np.zeros(0)) == 0.0


def _adversarial(kind, n):
    """Distributions that stress the radix select: everything in one round-0 bucket, the two
    middle ranks in different buckets, ties at the median, signed zeros, infinities, denormals."""
    rng = np.random.default_rng(len(kind) * 1000 + n)
    if kind == "equal":
        return np.full(n, 3.25)
    if kind == "zeros":
        x = np.zeros(n)
        x[::2] = -0.0
        return x
    if kind == "two":
        return rng.choice([-1.5, 2.5], size=n)
    if kind == "split":                      # lower half negative, upper half positive: ranks straddle the sign
        x = np.abs(rng.normal(size=n)) + 1.0
        x[: n // 2] *= -1.0
        return rng.permutation(x)
    if kind == "narrow":                     # 1 + k ulp: only the lowest bits differ
        return 1.0 + rng.integers(0, 1000, n) * np.finfo(np.float64).eps
    if kind == "denormal":
        return rng.integers(-50, 50, n).astype(np.float64) * 5e-324
    if kind == "inf":
        x = rng.normal(size=n)
        x[rng.integers(0, n, max(n // 3, 1))] = np.inf
        x[rng.integers(0, n, max(n // 3, 1))] = -np.inf
        return x
    if kind == "huge_range":
        return rng.normal(size=n) * 10.0 ** rng.integers(-300, 300, n)
    raise ValueError(kind)


@pytest.mark.parametrize("kind", ["equal", "zeros", "two", "split", "narrow", "denormal", "inf", "huge_range"])
@pytest.mark.parametrize("n", [2, 5, 64, 1000, 4097, 300_001])
def test_median_adversarial_bit_exact(cf, kind, n):
    x = _adversarial(kind, n)
    s = np.sort(x)
    want = s[n // 2] if n % 2 else 0.5 * (s[n // 2 - 1] + s[n // 2])     # median_dbl :280-292
    a = x.copy()
    got = cf.median(a)                  # srt = .true.: select on the device, then x is left sorted
    assert got == want or (np.isnan(got) and np.isnan(want)), (kind, n, got, want)
    assert np.array_equal(a, s)


def test_median_device_is_stream_ordered(cf):
    """cf_median_device with stream = NULL runs on the legacy default stream (like
    cf_moving_average_device), so it is ordered after work queued there."""
    import ctypes as C

    import torch
    from curvefit_b200 import _lib
    lib = _lib.load()
    n = 2_000_001
    x = torch.randn(n, dtype=torch.float64, device="cuda")
    y = x * 2.0 + 1.0                          # queued on the default stream right before the call
    out = torch.empty(1, dtype=torch.float64, device="cuda")
    rc = lib.cf_median_device(C.c_void_p(y.data_ptr()), n, C.c_void_p(out.data_ptr()), None)
    assert rc == 0
    assert float(out.item()) == float(torch.median(y).item())
