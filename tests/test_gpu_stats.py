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
