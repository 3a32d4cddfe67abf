"""GPU parity of the delta > 0 mode (src/curvefit_regression.f90:674-690)
against the oracle's lowess(delta): the set of fitted points is bit-exact, values within
1e-10.  The reference's smooth() fixes delta = 0 (:807), so the oracle is the only
statement of this mode's behaviour."""
import numpy as np
import pytest

from tests import cases

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def cf():
    from curvefit_b200 import regression
    return regression


def _delta_for(x, frac):
    return frac * float(x[-1] - x[0]) if x[-1] > x[0] else frac


@pytest.mark.parametrize("method", ["direct", "tree"])
@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal", "wide"])
@pytest.mark.parametrize("n,f,nsteps,frac", [(100, 0.3, 2, 0.05), (1000, 0.1, 3, 0.01), (4099, 0.3, 2, 0.02),
                                             (3000, 0.2, 0, 0.3), (2048, 0.05, 2, 1e-4), (513, 1.0, 1, 2.0),
                                             (20000, 0.05, 2, 0.005)])
def test_delta_matches_oracle(cf, oracle, method, kind, n, f, nsteps, frac):
    x = cases.make_x(kind, n, n + nsteps)
    y = cases.make_y(x, n + nsteps)
    delta = _delta_for(x, frac)
    ref = oracle.lowess(x, y, f, nsteps, delta=delta, want_windows=True)
    fit = cf.lowess_smoothing()
    fit.set_method(method)
    fit.initialize(x, y, srt=False)
    fit.set_delta(delta)
    ys = fit.smooth(f, nsteps=nsteps)
    # the points lowest() ran at: exactly the reference's
    fitted_ref = np.nonzero(ref["nleft"])[0] + 1
    assert np.array_equal(fit.fitted_points(), fitted_ref)
    scale = float(np.max(np.abs(ref["ys"])))
    d = float(np.max(np.abs(ys - ref["ys"])))
    assert d <= TOL * max(scale, 1e-300), ("ys", d, scale)
    d = float(np.max(np.abs(fit.m_residuals - ref["res"])))
    assert d <= TOL * max(float(np.max(np.abs(y))), 1e-300), ("res", d)
    if nsteps > 0:
        assert np.max(np.abs(fit.m_weights - ref["rw"])) <= 1e-9, "rw"


def test_near_ties_chain(cf, oracle):
    """Points closer than epsilon to their predecessor copy the fitted value (:685-688) --
    also when they are not bitwise equal -- but only while they stay within delta."""
    rng = np.random.default_rng(5)
    base = np.sort(rng.random(300)) + 0.5            # spacing of doubles in [0.5, 1): 1.1e-16 < eps
    x = np.sort(np.concatenate([base, np.nextafter(base[::3], 2.0), np.nextafter(np.nextafter(base[::6], 2.0), 2.0)]))
    y = cases.make_y(x, 7)
    for delta in (1e-17, 1e-16, 3e-16, 0.01):
        ref = oracle.lowess(x, y, 0.2, 2, delta=delta, want_windows=True)
        ys, rw, res = cf.cf_lowess_smooth_delta(x, y, 0.2, 2, delta)
        assert np.max(np.abs(ys - ref["ys"])) <= TOL * np.max(np.abs(ref["ys"])), delta
        assert np.max(np.abs(rw - ref["rw"])) <= 1e-9


def test_delta_zero_is_the_default_path(cf):
    x = cases.make_x("uniform", 5000, 1)
    y = cases.make_y(x, 1)
    a = cf.cf_lowess_smooth(x, y, 0.1, 2)
    b = cf.cf_lowess_smooth_delta(x, y, 0.1, 2, 0.0)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)


def test_delta_rejects_bad_input(cf):
    fit = cf.lowess_smoothing()
    x = cases.make_x("grid", 64)
    fit.initialize(x, cases.make_y(x))
    with pytest.raises(RuntimeError):
        fit.set_delta(-1.0)
    with pytest.raises(RuntimeError):
        fit.set_delta(float("nan"))


def test_large_n_delta(cf, oracle):
    """n = 2e6, delta = 0.2 % of the range: ~500 fitted points; the oracle runs it in full
    because only they cost O(ns)."""
    from curvefit_b200 import datagen
    n, f = 2_000_000, 0.01
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x, 11)
    delta = 0.002
    ref = oracle.lowess(x, y, f, 2, delta=delta)
    for method in ("direct", "tree"):
        fit = cf.lowess_smoothing()
        fit.set_method(method)
        fit.initialize(x, y, srt=False)
        fit.set_delta(delta)
        ys = fit.smooth(f, nsteps=2)
        assert fit.fitted_points().size < 1000
        assert np.max(np.abs(ys - ref["ys"])) <= TOL * np.max(np.abs(ref["ys"])), method
        assert np.max(np.abs(fit.m_weights - ref["rw"])) <= 1e-9


def test_delta_refused_on_a_span_upload(cf):
    """After cf_lowess_init_span x holds -inf / +inf sentinels outside the uploaded span, so the
    fitted-point list of delta > 0 (:682-690) cannot be derived from it: both orders of the two
    calls are refused with CF_INVALID_INPUT_ERROR, and a full init lifts the refusal."""
    import ctypes as C

    from curvefit_b200 import CF_INVALID_INPUT_ERROR, _lib
    lib = _lib.load()
    n = 20_000
    x = cases.make_x("grid", n)
    y = cases.make_y(x)
    xp, yp = x.ctypes.data_as(_lib.c_dp), y.ctypes.data_as(_lib.c_dp)
    h = C.c_void_p()
    assert lib.cf_lowess_create(C.byref(h), -1, None) == 0
    try:
        # span upload first, delta second
        assert lib.cf_lowess_init_span(h, xp, yp, n, 5_000, 12_000) == 0
        assert lib.cf_lowess_set_delta(h, 0.01) == CF_INVALID_INPUT_ERROR
        buf = C.create_string_buffer(256)
        lib.cf_last_error(buf, 256)
        assert b"whole of x" in buf.value
        assert lib.cf_lowess_delta(h) == 0.0
        # delta first (on a fully initialised handle), span upload second: the run refuses
        assert lib.cf_lowess_init(h, xp, yp, n, 0) == 0
        assert lib.cf_lowess_set_delta(h, 0.01) == 0
        assert lib.cf_lowess_run(h, 0.1, 1, None) == 0
        assert lib.cf_lowess_init_span(h, xp, yp, n, 5_000, 12_000) == 0
        assert lib.cf_lowess_run(h, 0.1, 1, None) == CF_INVALID_INPUT_ERROR
        # a span that covers everything writes no sentinels and is accepted
        assert lib.cf_lowess_init_span(h, xp, yp, n, 1, n) == 0
        assert lib.cf_lowess_run(h, 0.1, 1, None) == 0
        # delta = 0 on a span upload is unaffected
        assert lib.cf_lowess_set_delta(h, 0.0) == 0
        assert lib.cf_lowess_init_span(h, xp, yp, n, 5_000, 12_000) == 0
    finally:
        lib.cf_lowess_destroy(h)
