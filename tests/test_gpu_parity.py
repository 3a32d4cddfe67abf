"""GPU parity tests: the CUDA path (through the C ABI) against the oracle on the
same seeded inputs.  Indices bit-exact; values within 1e-10 (north_star's
tolerance, with the floor SURVEY 7.4-2 defines: |d_i| <= 1e-10 *
max(|ref_i|, 1e-3 max|ref|) and normwise 1e-10)."""
import os

import numpy as np
import pytest

from curvefit_b200 import datagen
from tests import cases

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10


@pytest.fixture(scope="module")
def cf():
    import curvefit_b200
    from curvefit_b200 import regression
    return regression


def _check_windows(fit, oracle, x, y, f):
    """nleft/nright equal the slide loop's; [lo9, hi9] equals the reference's
    contributing set {j >= nleft : r <= h9}, whose last member is nrt or nrt-1
    (the sweep stops ON the first excluded point to the right, :553-554)."""
    r = oracle.lowess(x, y, f, 0, want_windows=True)
    w = fit.windows(f)
    lead = np.nonzero(r["nleft"])[0]
    assert np.array_equal(w["leader"][lead], lead + 1)
    assert np.array_equal(w["nleft"][lead], r["nleft"][lead])
    assert np.array_equal(w["nright"][lead], r["nright"][lead])
    n = x.size
    for i in lead[:: max(1, lead.size // 400)]:
        nl, nr = r["nleft"][i] - 1, r["nright"][i] - 1
        h = max(x[i] - x[nl], x[nr] - x[i])
        rr = np.abs(x - x[i])
        inside = np.nonzero((rr <= 0.999 * h) & (np.arange(n) >= nl))[0]
        assert w["lo9"][i] - 1 == inside[0] and w["hi9"][i] - 1 == inside[-1], i
        core = np.nonzero((rr <= 0.001 * h) & (np.arange(n) >= nl))[0]
        assert w["lo1"][i] - 1 == core[0] and w["hi1"][i] - 1 == core[-1], i
        nrt = r["nrt"][i] - 1
        assert inside[-1] in (nrt, nrt - 1) or nrt == n - 1
    foll = np.nonzero(r["nleft"] == 0)[0]
    if foll.size:
        assert np.all(x[w["leader"][foll] - 1] == x[foll])
        assert np.all(w["leader"][foll] - 1 < foll)


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal"])
@pytest.mark.parametrize("n,f", [(2, 0.5), (3, 1.0), (100, 0.2), (100, 0.8), (1000, 0.05),
                                 (4099, 0.3), (16384, 0.3), (5000, 1.0), (777, 0.0)])
def test_windows_bit_exact(cf, oracle, kind, n, f):
    x = cases.make_x(kind, n, n)
    y = cases.make_y(x, n)
    fit = cf.lowess_smoothing()
    fit.initialize(x, y, srt=False)
    _check_windows(fit, oracle, x, y, f)


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal", "wide"])
# (n = 3, f = 1 interpolates the end points exactly: its residuals are rounding noise, so
# a robust iteration on them is chaotic in ANY implementation -- checked with nsteps = 0)
@pytest.mark.parametrize("n,f,nsteps", [(2, 0.5, 2), (3, 1.0, 0), (100, 0.2, 2), (100, 0.8, 2),
                                        (1000, 0.05, 3), (4099, 0.3, 2), (513, 1.0, 1),
                                        (2048, 0.01, 2), (600, 0.4, 0), (3000, 0.2, 5),
                                        (16384, 0.3, 2)])
def test_smooth_matches_oracle(cf, oracle, kind, n, f, nsteps):
    x = cases.make_x(kind, n, n + nsteps)
    y = cases.make_y(x, n + nsteps)
    ref = oracle.lowess_mt(x, y, f, nsteps) if n > 5000 else oracle.lowess(x, y, f, nsteps)
    fit = cf.lowess_smoothing()
    fit.set_method("direct")
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=nsteps)
    # pointwise 1e-10 against the oracle; points where the oracle's OWN rounding exceeds
    # that (zero crossings at n >= 16384) are refereed in extended precision
    ok, d, nref = cases.close_or_exact(ys, ref["ys"], x, y, fit.m_weights if nsteps > 0 else None, f, TOL)
    if x.size < 16384:
        assert nref == 0, f"{nref} point(s) needed the referee at n = {x.size}"
    assert ok, ("ys", d, nref)
    # res = y - ys carries the ABSOLUTE error of ys, so its scale is y's, not its own
    d = float(np.max(np.abs(fit.m_residuals - ref["res"])))
    assert d <= TOL * max(float(np.max(np.abs(y))), 1e-300), ("res", d)
    if nsteps > 0:
        assert np.max(np.abs(fit.m_weights - ref["rw"])) <= 1e-9, "rw"
        assert abs(fit.last_stats.cmad_last - ref["cmad"][-1]) <= 1e-10 * max(ref["cmad"][-1], 1e-300)


def test_example2_through_reference_style_api(cf, oracle):
    """tests/curvefit_test_regression.f90:26-38: initialize, smooth(0.2), smooth(0.8)
    on one object; checked against the committed golden vectors."""
    g = np.load(os.path.join(GOLD, "c1_example2.npz"))
    x, y = datagen.example2(100)
    fit = cf.lowess_smoothing()
    fit.initialize(x, y)
    assert fit.get_count() == 100 and fit.get_x(1) == x[0] and fit.get_y(100) == y[99]
    ys = fit.smooth(0.2)
    ys2 = fit.smooth(0.8)
    assert cases.close(ys, g["f02_ys"], TOL)[0]
    assert cases.close(ys2, g["f08_ys"], TOL)[0]
    buf = np.zeros(40)
    fit.get_residuals(buf)
    assert cases.close(buf, g["f08_res"][:40], TOL)[0]
    assert np.max(np.abs(fit.m_weights - g["f08_rw"])) < 1e-9


@pytest.mark.parametrize("name", ["n4096_uniform", "n1000_ties"])
def test_golden_vectors(cf, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    ys, rw, res = cf.cf_lowess_smooth(g["x"], g["y"], float(g["f"]), int(g["nsteps"]))
    assert cases.close(ys, g["ys"], TOL)[0]
    assert cases.close(res, g["res"], TOL)[0]
    assert np.max(np.abs(rw - g["rw"])) < 1e-9


def test_select_and_weights_bit_exact_given_same_residuals(cf, oracle):
    """F6: the radix select equals the reference's sort-and-index exactly, and the
    bisquare weights are bit-identical given the same residuals."""
    rng = np.random.default_rng(3)
    for n in (2, 3, 10, 1001, 4096, 50001):
        x = np.sort(rng.random(n))
        y = rng.normal(size=n)
        fit = cf.lowess_smoothing()
        fit.initialize(x, y, srt=False)
        ys, rw, res = fit.run_from_rw(0.3, None, passes=2)   # pass, robust update, pass
        # res is of pass 2; recompute pass-1 residuals to check the select input
        ys1, _, res1 = fit.run_from_rw(0.3, None, passes=1)
        a = np.sort(np.abs(res1))
        m1, m2 = n // 2, n - (1 + n // 2)
        cmad = 3.0 * (a[m1] + a[m2])
        c9, c1 = 0.999 * cmad, 0.001 * cmad
        r = np.abs(res1)
        with np.errstate(invalid="ignore", divide="ignore"):
            t = r / cmad
            want = np.where(r <= c1, 1.0, np.where(r > c9, 0.0, (1.0 - t * t) * (1.0 - t * t)))
        _, rw2, _ = fit.run_from_rw(0.3, None, passes=2)
        assert np.array_equal(rw2, want), n


def test_zero_mad_and_dead_windows(cf, oracle):
    """More than half of the residuals exactly zero -> cmad = 0 -> rw in {0, 1}
    (:706-711); windows whose robustness weights are all zero -> ok = .false. ->
    ys = y (:559, :673).  y is zero except a block of non-collinear outliers, so
    the zero residuals are exact zeros in any arithmetic."""
    n = 60
    x = datagen.grid_x(n)
    y = np.zeros(n)
    y[20:26] = [5.0, -3.0, 7.0, -4.0, 6.0, -2.0]
    ref = oracle.lowess(x, y, 0.1, 2)
    assert ref["cmad"][-1] == 0.0 and set(np.unique(ref["rw"])) <= {0.0, 1.0}
    ys, rw, res = cf.cf_lowess_smooth(x, y, 0.1, 2)
    assert np.array_equal(rw, ref["rw"])
    assert cases.close(ys, ref["ys"], TOL)[0]
    dead = np.nonzero(ref["ys"] == y)[0]
    assert np.array_equal(ys[dead], y[dead])


def test_sort_path_matches_ls_init(cf, oracle):
    rng = np.random.default_rng(8)
    for n in (1, 2, 1000, 4096, 4097, 100003):
        x = rng.normal(size=n) * 10.0
        y = rng.normal(size=n)
        xs, ys, perm = cf.cf_sort_xy(x, y)
        o = np.argsort(x, kind="stable")
        assert np.array_equal(xs, x[o]) and np.array_equal(ys, y[o])
        assert np.array_equal(perm, o + 1)
    # ties: sorted x identical to the reference's; y equal per tie group as multisets
    x = np.round(rng.random(5000) * 50) / 50
    y = rng.normal(size=5000)
    xs, ys, _ = cf.cf_sort_xy(x, y)
    mx, my = oracle.ls_init(x, y, True)
    assert np.array_equal(xs, mx)
    for v in np.unique(xs):
        assert np.array_equal(np.sort(ys[xs == v]), np.sort(my[mx == v]))
    # negative zero, denormals, infinities keep IEEE order
    x = np.array([3.0, -0.0, 0.0, -1e-310, 1e-310, -np.inf, np.inf, -2.5])
    xs, _, _ = cf.cf_sort_xy(x, x.copy())
    assert np.array_equal(xs, np.sort(x))


def test_initialize_with_sort_then_smooth(cf, oracle):
    n = 20000
    x = datagen.random_x(n, 1)
    y = datagen.noisy_y(x, 1)
    fit = cf.lowess_smoothing()
    fit.initialize(x, y)                      # srt defaults to .true. (:760)
    o = np.argsort(x, kind="stable")
    assert np.array_equal(fit.m_x, x[o]) and np.array_equal(fit.m_y, y[o])
    ys = fit.smooth(0.05)
    ref = oracle.lowess_mt(x[o], y[o], 0.05, 2)
    assert cases.close(ys, ref["ys"], TOL)[0]


def test_batched_matches_single(cf, oracle):
    s, n = 12, 2048
    x = np.stack([np.sort(datagen.random_x(n, k)) for k in range(s)])
    y = np.stack([datagen.noisy_y(x[k], k) for k in range(s)])
    ys, rw, res = cf.cf_lowess_smooth_batched(x, y, 0.3, 2)
    for k in range(s):
        ref = oracle.lowess(x[k], y[k], 0.3, 2)
        assert cases.close(ys[k], ref["ys"], TOL)[0], k
        assert cases.close(res[k], ref["res"], TOL)[0], k


def test_batched_resident_handle_and_large_group(cf, oracle):
    """Series as a grid dimension: 300 series of n = 700 (K-way split active), checked
    in full for a few series; the resident handle gives the same bits as the
    host-buffer entry point."""
    s, n, f = 300, 700, 0.25
    x = np.stack([np.sort(datagen.random_x(n, k)) for k in range(s)])
    y = np.stack([datagen.noisy_y(x[k], k) for k in range(s)])
    b = cf.batched_lowess()
    b.initialize(x, y)
    b.run_resident(f, 3)
    ys, rw, res = b.fetch()
    ys2, rw2, res2 = cf.cf_lowess_smooth_batched(x, y, f, 3)
    assert np.array_equal(ys, ys2) and np.array_equal(rw, rw2) and np.array_equal(res, res2)
    for k in (0, 1, 137, 299):
        ref = oracle.lowess(x[k], y[k], f, 3)
        assert cases.close(ys[k], ref["ys"], TOL)[0], k
        assert np.max(np.abs(rw[k] - ref["rw"])) < 1e-9, k


def test_shards_reproduce_the_single_gpu_result_bitwise(cf):
    """8e determinism: fitting block g of G gives bit-identical ys on that block."""
    import ctypes as C
    from curvefit_b200 import _lib
    n, f = 6000, 0.1
    x = np.sort(datagen.random_x(n, 2))
    y = datagen.noisy_y(x, 2)
    whole = cf.lowess_smoothing()
    whole.initialize(x, y, srt=False)
    ys_all, _, _ = whole.run_from_rw(f, None, 1)
    lib = _lib.load()
    for G in (2, 3):
        for g in range(G):
            part = cf.lowess_smoothing()
            part.initialize(x, y, srt=False)
            assert lib.cf_lowess_set_shard(part._h, g, G) == 0
            b, e, blk = C.c_int32(), C.c_int32(), C.c_int32()
            lib.cf_lowess_shard(part._h, C.byref(b), C.byref(e), C.byref(blk))
            ys, _, _ = part.run_from_rw(f, None, 1)
            assert np.array_equal(ys[b.value:e.value], ys_all[b.value:e.value]), (G, g)


def test_large_series_teacher_forced_samples(cf, oracle):
    """n = 2e5, f = 0.8, nsteps = 5 (scaled-down C5): full-size GPU run, oracle
    evaluated at sampled points with the GPU's own robustness weights (8c)."""
    n, f = 200000, 0.8
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    fit = cf.lowess_smoothing()
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=5)
    assert np.all(np.isfinite(ys))
    rw_in = fit.m_weights.copy()
    ys_next, _, _ = fit.run_from_rw(f, rw_in, passes=1)
    idx = np.unique(np.concatenate([np.arange(0, 3), np.arange(n - 3, n),
                                    np.linspace(0, n - 1, 120).astype(np.int64)])).astype(np.int32)
    t = oracle.lowest_at(x, y, f, idx, rw_in)
    ok, d, nref = cases.close_or_exact(ys_next[idx], t["ys"], x, y, rw_in, f, TOL, idx=idx)
    assert ok, (d, nref)


@pytest.mark.parametrize("method", ["direct", "tree"])
@pytest.mark.parametrize("nsteps,delta", sorted(cases.NETLIB_YS))
def test_published_netlib_example_through_the_c_abi(cf, oracle, method, nsteps, delta):
    """The worked example in the header of netlib's lowess.f (tests/cases.NETLIB_*: ten tied
    abscissas, delta interpolation, robustness iterations): both evaluators print the
    published three decimals and agree with the oracle to 1e-10."""
    x, y, f = cases.NETLIB_X, cases.NETLIB_Y, cases.NETLIB_F
    want = np.array(cases.NETLIB_YS[(nsteps, delta)])
    fit = cf.lowess_smoothing()
    fit.set_method(method)
    fit.initialize(x, y, srt=False)
    if delta > 0.0:
        fit.set_delta(delta)
    ys = fit.smooth(f, nsteps=nsteps)
    assert np.array_equal(np.round(ys, 3), want), np.round(ys, 3) - want
    ref = oracle.lowess(x, y, f, nsteps, delta)
    assert cases.close(ys, ref["ys"], 1e-10)[0]
    # the stateless host-buffer entry points the Fortran shim binds
    if delta > 0.0:
        ys2, _, res2 = cf.cf_lowess_smooth_delta(x, y, f, nsteps, delta)
    else:
        ys2, _, res2 = cf.cf_lowess_smooth(x, y, f, nsteps)
    assert np.array_equal(np.round(ys2, 3), want)
    assert np.array_equal(res2, y - ys2)


@pytest.mark.parametrize("kind", ["exact_line", "one_outlier", "quantised", "tiny"])
def test_robust_step_on_degenerate_residuals(cf, oracle, kind):
    """The select behind the robustness step (:697-703) where all |res| share one bucket (exact
    fit: res = 0 or a few ulp), where the two middle ranks differ, and on heavy ties."""
    n, f, nsteps = 6001, 0.3, 2
    x = cases.make_x("uniform", n, 7)
    rng = np.random.default_rng(11)
    if kind == "exact_line":
        y = 2.0 * x + 1.0
    elif kind == "one_outlier":
        y = 2.0 * x + 1.0
        y[n // 3] += 5.0
    elif kind == "quantised":
        y = np.round(cases.make_y(x, 7) * 4.0) / 4.0
    else:
        y = (2.0 * x + 1.0) + rng.choice([-1e-300, 0.0, 1e-300], n)
    ref = oracle.lowess(x, y, f, nsteps)
    ys, rw, res = cf.cf_lowess_smooth(x, y, f, nsteps)
    assert cases.close(ys, ref["ys"], TOL)[0]
    # the weights follow from (res, cmad) exactly: recompute the reference step on OUR residuals
    fit = cf.lowess_smoothing()
    fit.initialize(x, y, srt=False)
    ys1, rw1, res1 = fit.run_from_rw(f, None, passes=2)
    ys0, _, res0 = fit.run_from_rw(f, None, passes=1)
    cmad, rw_ref = oracle.robust_weights(res0)
    assert np.array_equal(rw1, rw_ref), kind


def test_fewer_than_two_points_return_y_and_touch_nothing_else(cf):
    """lowess with n < 2 (:651-654): ys = y and the routine returns before rw or res are
    written; n = 0 is a no-op.  Through the stateless call, the multi-GPU call (one device
    here) and the type."""
    from curvefit_b200 import _lib
    lib = _lib.load()
    p = lambda a: a.ctypes.data_as(_lib.c_dp)
    for fn, extra in ((lib.cf_lowess_smooth, ()), (lib.cf_lowess_smooth_multi, (None, 0))):
        x, y = np.array([0.25]), np.array([3.5])
        ys, rw, res = np.full(1, -7.0), np.full(1, -8.0), np.full(1, -9.0)
        assert fn(p(x), p(y), 1, 0.5, 2, *extra, p(ys), p(rw), p(res)) == 0
        assert ys[0] == 3.5 and rw[0] == -8.0 and res[0] == -9.0
        e = np.zeros(0)
        assert fn(p(e), p(e), 0, 0.5, 2, *extra, p(e), p(e), p(e)) == 0
        assert fn(p(x), p(y), -1, 0.5, 2, *extra, p(ys), p(rw), p(res)) == 202      # CF_ARRAY_SIZE_ERROR
        assert fn(p(x), p(y), 1, 0.5, -1, *extra, p(ys), p(rw), p(res)) == 201     # CF_INVALID_INPUT_ERROR
    fit = cf.lowess_smoothing()
    fit.initialize(np.array([0.25]), np.array([3.5]))
    assert fit.get_count() == 1
    assert np.array_equal(fit.smooth(0.3), np.array([3.5]))


def test_per_call_evaluator_from_two_threads():
    """cf_lowess_smooth_ex: the evaluator is an argument of the call, local to the calling thread.
    Two threads smooth the same series at the same time, one with the direct sweep and one with the
    moment tree; each must get exactly what a handle set to its evaluator gives, and the process
    default must not move."""
    import threading

    from curvefit_b200 import _lib, lowess_smoothing
    lib = _lib.load()
    n, f, nsteps = 300_000, 0.05, 2
    x = cases.make_x("uniform", n, 11)
    y = cases.make_y(x, 11)
    want = {}
    for name in ("direct", "tree"):
        fit = lowess_smoothing()
        fit.set_method(name)
        fit.initialize(x, y, srt=False)
        want[name] = fit.smooth(f, nsteps=nsteps).copy()
        fit.close()
    assert not np.array_equal(want["direct"], want["tree"])       # else the test could not tell them apart
    p = lambda a: a.ctypes.data_as(_lib.c_dp)
    default_before = lib.cf_default_method()
    bad = []

    def work(name, method):
        ys, rw, res = np.empty(n), np.empty(n), np.empty(n)
        for _ in range(6):
            rc = lib.cf_lowess_smooth_ex(p(x), p(y), n, f, nsteps, method, p(ys), p(rw), p(res))
            if rc != 0 or not np.array_equal(ys, want[name]):
                bad.append((name, rc))

    th = [threading.Thread(target=work, args=("direct", _lib.CF_METHOD_DIRECT)),
          threading.Thread(target=work, args=("tree", _lib.CF_METHOD_TREE))]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not bad, bad
    assert lib.cf_default_method() == default_before
    out = np.empty(n)
    assert lib.cf_lowess_smooth_ex(p(x), p(y), n, f, nsteps, 7, p(out), p(out), p(out)) == 201
