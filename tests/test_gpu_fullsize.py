"""GPU parity at BASELINE.json's FULL sizes (C2, C3, C4, C5), where the oracle cannot run
a whole smooth() (SURVEY 8c "oracle at sizes it cannot run end-to-end"):

  * teacher forcing -- the oracle's lowest() at stratified sample points (both ends,
    interior), fed with the robustness weights the GPU itself entered the pass with;
  * the robustness step in full, O(n) on the host: cmad = 3 * (the two middle order
    statistics of |res|) by numpy partition, bisquare weights recomputed from (res, cmad)
    -- both BIT-exact (src/curvefit_regression.f90:697-713 has no summation order);
  * size-independent properties: exactly linear data is reproduced (ys = y to 1e-10,
    every robustness weight 1); res == y - ys bitwise.
Tolerance for values: 1e-10 (tests/cases.close_or_exact)."""
import numpy as np
import pytest

from curvefit_b200 import datagen
from tests import cases

pytestmark = pytest.mark.gpu
TOL = 1e-10


@pytest.fixture(scope="module")
def cf():
    from curvefit_b200 import regression
    return regression


def _sample_idx(n, m):
    return np.unique(np.concatenate([np.arange(0, 3), np.arange(n - 3, n),
                                     np.linspace(0, n - 1, m).astype(np.int64)])).astype(np.int32)


def _bisquare(res, cmad):
    """:704-713 with the reference's operation order (numpy float64 = IEEE, no FMA)."""
    r = np.abs(res)
    c9, c1 = 0.999 * cmad, 0.001 * cmad
    with np.errstate(divide="ignore", invalid="ignore"):
        t = r / cmad
        u = 1.0 - t * t
        w = u * u
    w = np.where(r <= c1, 1.0, np.where(r > c9, 0.0, w))
    return w


def _check_robust_step(fit, f, y):
    """One more pass from the GPU's own weights, then the robustness step on its residuals,
    verified in full on the host."""
    n = y.size
    rw_in = fit.m_weights.copy()
    ys1, rw1, res1 = fit.run_from_rw(f, rw_in, passes=2)      # pass, robust update, pass
    ys0, _, res0 = fit.run_from_rw(f, rw_in, passes=1)        # the first of the two passes again
    assert np.array_equal(res0, y - ys0)                      # :695, bitwise
    a = np.abs(res0)
    m1, m2 = n // 2, n - (1 + n // 2)                         # 0-based ranks of rw(m1), rw(m2), :699-700
    part = np.partition(a, [m2, m1])
    cmad = 3.0 * (part[m1] + part[m2])                        # :701
    assert np.array_equal(rw1, _bisquare(res0, cmad))         # select + bisquare: bit-exact
    return rw_in, ys0


def _teacher(cf_fit, oracle, x, y, f, rw_in, ys_pass, m):
    idx = _sample_idx(x.size, m)
    t = oracle.lowest_at(x, y, f, idx, rw_in)
    ok, d, nref = cases.close_or_exact(ys_pass[idx], t["ys"], x, y, rw_in, f, TOL, idx=idx)
    assert ok, (d, nref)


def test_c2_full_direct_kernel(cf, oracle):
    """BASELINE configs[1]: n = 1e6, f = 0.05, 3 robustness iterations, direct kernel."""
    n, f, nsteps = 1_000_000, 0.05, 3
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    fit = cf.lowess_smoothing()
    fit.set_method("direct")
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=nsteps)
    assert np.all(np.isfinite(ys)) and np.array_equal(fit.m_residuals, y - ys)
    rw_in, ys0 = _check_robust_step(fit, f, y)
    _teacher(fit, oracle, x, y, f, rw_in, ys0, 200)


@pytest.mark.parametrize("name,n,f,nsteps", [("C2 (BASELINE configs[1])", 1_000_000, 0.05, 3),
                                             ("C5 scaled (f = 0.8, 5 iterations)", 200_000, 0.8, 5)])
def test_every_point_against_a_complete_oracle_run(cf, oracle, name, n, f, nsteps):
    """No teacher forcing: the oracle smooths the WHOLE series on all host cores (lowess_mt, about
    a minute each) and every one of the n values of ys, rw and res after all nsteps + 1 passes is
    compared, for both evaluators -- what a user switching from the reference would see."""
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    oracle.set_threads(oracle.host_cores())
    ref = oracle.lowess_mt(x, y, f, nsteps)
    scale = float(np.max(np.abs(ref["ys"])))
    for method in ("direct", "tree"):
        fit = cf.lowess_smoothing()
        fit.set_method(method)
        fit.initialize(x, y, srt=False)
        ys = fit.smooth(f, nsteps=nsteps)
        d_ys = float(np.max(np.abs(ys - ref["ys"])))
        d_res = float(np.max(np.abs(fit.m_residuals - ref["res"])))
        d_rw = float(np.max(np.abs(fit.m_weights - ref["rw"])))
        assert d_ys <= TOL * scale, (name, method, d_ys)
        assert d_res <= TOL * scale, (name, method, d_res)
        # a weight is a smooth function of res / cmad (slope <= 2 / cmad, cmad ~ 1.5 here) except
        # for the 4e-6 step at r = .999 cmad (:707-710): nothing sits that close in these series
        assert d_rw <= 1e-9, (name, method, d_rw)
        fit.close()


def test_c4_full_tree_and_linear_property(cf, oracle):
    """BASELINE headline series: n = 1e8, f = 0.05, 3 robustness iterations (moment-tree
    evaluator; the direct kernel needs > 15 min per smooth at this size)."""
    n, f, nsteps = 100_000_000, 0.05, 3
    x = datagen.grid_x(n)
    # linearity first (cheap to state, exact to check): y = a + b x  ->  ys = y, rw = 1
    y = 2.0 - 3.0 * x
    fit = cf.lowess_smoothing()
    fit.set_method("tree")
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=1)
    assert np.max(np.abs(ys - y)) <= TOL * np.max(np.abs(y))
    del ys
    y = datagen.noisy_y(x)
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=nsteps)
    assert np.all(np.isfinite(ys)) and np.array_equal(fit.m_residuals, y - ys)
    del ys
    rw_in, ys0 = _check_robust_step(fit, f, y)
    _teacher(fit, oracle, x, y, f, rw_in, ys0, 36)


def test_c4_direct_kernel_sampled_pass(cf, oracle):
    """The DIRECT kernel at the headline geometry n = 1e8, ns = 5e6 (a full pass is 5.5 min):
    one weighted pass over a strided sample of one wave of 512-point tiles plus the first and
    the last tile, each evaluated exactly as in a full pass (K = 1, 19 532 staged chunks per
    tile), against the oracle's lowest() fed with the same robustness weights, and against the
    moment tree's pass from the same weights."""
    n, f = 100_000_000, 0.05
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    fit = cf.lowess_smoothing()
    fit.set_method("tree")
    fit.initialize(x, y, srt=False)
    fit.smooth(f, nsteps=1)                       # robustness weights of one iteration
    rw_in = fit.m_weights.copy()
    ys_tree, _, _ = fit.run_from_rw(f, rw_in, passes=1)
    tiles = -(-n // 512)
    stride = tiles // 592
    ms, idx, ysd = fit.fit_sample(f, True, stride // 2, stride, 592)
    assert idx.size == 592 * 512 and ms > 0
    ms2, idx2, ysd2 = fit.fit_sample(f, True, 0, tiles - 1, 2)      # both ends of the series
    idx = np.concatenate([idx, idx2]); ysd = np.concatenate([ysd, ysd2])
    assert np.all(np.isfinite(ysd))
    assert cases.close(ysd, ys_tree[idx], TOL)[0]
    pick = np.unique(np.concatenate([np.arange(0, 3), np.arange(idx.size - 3, idx.size),
                                     np.linspace(0, idx.size - 1, 26).astype(np.int64)]))
    t = oracle.lowest_at(x, y, f, idx[pick].astype(np.int32), rw_in)
    ok, d, nref = cases.close_or_exact(ysd[pick], t["ys"], x, y, rw_in, f, TOL, idx=idx[pick].astype(np.int32))
    assert ok and nref == 0, (d, nref)


def test_c5_full_wide_span_both_evaluators(cf, oracle):
    """BASELINE configs[4]: n = 2e6, f = 0.8, 5 robustness iterations.  The whole smooth runs
    on the tree; one weighted pass of the direct kernel (2 s) must agree with it and with
    the oracle's samples."""
    n, f, nsteps = 2_000_000, 0.8, 5
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    fit = cf.lowess_smoothing()
    fit.set_method("tree")
    fit.initialize(x, y, srt=False)
    ys = fit.smooth(f, nsteps=nsteps)
    assert np.all(np.isfinite(ys))
    rw_in, ys0 = _check_robust_step(fit, f, y)
    _teacher(fit, oracle, x, y, f, rw_in, ys0, 48)
    fit.set_method("direct")
    ysd, _, _ = fit.run_from_rw(f, rw_in, passes=1)
    assert cases.close(ysd, ys0, TOL)[0]
    _teacher(fit, oracle, x, y, f, rw_in, ysd, 48)


def test_c3_full_batched(cf, oracle):
    """BASELINE configs[2]: 4096 independent series x n = 16384, f = 0.3 (nsteps = 2), through
    the host-buffer batched entry point; four series are checked in full against the oracle,
    every series by the bitwise residual identity and the weight bounds."""
    s, n, f = 4096, 16384, 0.3
    x1 = datagen.grid_x(n)
    base = datagen.signal(x1)
    x = np.broadcast_to(x1, (s, n)).copy()
    y = np.empty((s, n))
    for k in range(s):
        y[k] = base + (datagen.uniform01(datagen.SEED_Y + (k << 32), n) - 0.5)
    ys, rw, res = cf.cf_lowess_smooth_batched(x, y, f, 2)
    assert np.all(np.isfinite(ys)) and np.array_equal(res, y - ys)
    assert np.all((rw >= 0.0) & (rw <= 1.0))
    for k in (0, 1, 2047, 4095):
        ref = oracle.lowess_mt(x[k], y[k], f, 2)
        ok, d, nref = cases.close_or_exact(ys[k], ref["ys"], x[k], y[k], rw[k], f, TOL)
        assert ok, (k, d, nref)
        assert np.max(np.abs(rw[k] - ref["rw"])) < 1e-9, k


@pytest.mark.parametrize("kind", ["uniform", "clustered", "ties", "wide"])
def test_windows_guided_searches_large_irregular(cf, kind):
    """windows.cu finds every index as "the first j of [lo, hi] where a monotone predicate
    holds" from a GUESS (carried from the thread's previous point when n is large enough for
    several points per thread -- here 4).  The defining property is checked for EVERY point
    with the reference's own floating-point expressions (numpy float64: IEEE, no FMA):
    the predicate holds at the answer (or the answer is the interval's assumed end) and
    fails just before it (or the answer is the interval's start).  For a monotone predicate
    that pins the index uniquely, so this is a bit-exact check at a size the oracle's
    O(n ns) sweep cannot reach.  (:664-666 slide predicate, :542-556 r <= h9 / r <= h1.)"""
    n, f = 2_500_000, 0.05
    x = cases.make_x(kind, n, 7)
    y = datagen.noisy_y(x)
    fit = cf.lowess_smoothing()
    fit.initialize(x, y, srt=False)
    w = fit.windows(f)
    from curvefit_b200 import _lib
    ns = int(_lib.load().cf_lowess_span(n, f))                        # :647
    assert ns == max(min(int(f * n), n), 2)
    i = w["leader"].astype(np.int64) - 1
    assert np.all(x[i] == x) and np.all(i <= np.arange(n))
    assert np.all((i == 0) | (x[np.maximum(i - 1, 0)] < x[i]))          # first of its tie group
    xs = x[i]
    nl = w["nleft"].astype(np.int64) - 1
    nr = w["nright"].astype(np.int64) - 1
    l9, r9 = w["lo9"].astype(np.int64) - 1, w["hi9"].astype(np.int64) - 1
    l1, r1 = w["lo1"].astype(np.int64) - 1, w["hi1"].astype(np.int64) - 1
    assert np.array_equal(nr, nl + ns - 1) and nl.min() >= 0 and nl.max() <= n - ns

    def slide(L):                                  # x(i) - x(L) <= x(L + ns) - x(i), L + ns < n
        Lc = np.clip(L, 0, n - ns - 1)
        return (xs - x[Lc]) <= (x[Lc + ns] - xs)
    assert np.all((nl == n - ns) | slide(nl))
    assert np.all((nl == 0) | ~slide(nl - 1))

    h = np.maximum(xs - x[nl], x[nr] - xs)
    h9, h1 = 0.999 * h, 0.001 * h

    def inside(j, hh):
        return np.abs(x[np.clip(j, 0, n - 1)] - xs) <= hh
    assert np.all((nl <= l9) & (l9 <= l1) & (l1 <= i) & (i <= r1) & (r1 <= r9) & (r9 <= n - 1))
    assert np.all((l9 == i) | inside(l9, h9)) and np.all((l9 == nl) | ~inside(l9 - 1, h9))
    assert np.all((r9 == i) | inside(r9, h9)) and np.all((r9 == n - 1) | ~inside(r9 + 1, h9))
    assert np.all((l1 == i) | inside(l1, h1)) and np.all((l1 == l9) | ~inside(l1 - 1, h1))
    assert np.all((r1 == i) | inside(r1, h1)) and np.all((r1 == r9) | ~inside(r1 + 1, h1))


@pytest.mark.parametrize("kind", ["grid", "ties", "run"])
def test_host_buffer_smooth_with_overlapped_copies_equals_resident_run(cf, kind):
    """cf_lowess_smooth on a long series overlaps its copies with the kernels: direct evaluator --
    y's upload beside the window search, the final weights' download beside the last pass; tree
    evaluator -- the FIRST pass (windows, tree build, evaluation) follows the chunked upload, and
    ys / res leave piece by piece during the last pass, which is evaluated in pieces.  Every output
    must equal the plain sequence init -> run -> fetch bit for bit."""
    n, f, nsteps = 6_000_000, 0.002, 2
    if kind == "grid":
        x = datagen.grid_x(n)
    elif kind == "ties":
        x = np.sort(np.round(datagen.random_x(n, 9) * (n // 6)) / (n // 6))
    else:
        # a run of 1.5e6 equal abscissas across three upload chunks: windows inside it have h = 0 and
        # take the whole run (:546) -- its points may only be processed once the run has ended
        u = datagen.random_x(n - 1_500_000, 9)
        x = np.sort(np.concatenate([0.4 * u[: u.size // 2], np.full(1_500_000, 0.5), 0.6 + 0.4 * u[u.size // 2:]]))
    y = datagen.noisy_y(x, 9)
    plan = (("tree", 2), ("tree", 0)) if kind == "run" else (("tree", 2), ("direct", 2), ("tree", 0))
    for method, nsteps in plan:
        fit = cf.lowess_smoothing()
        fit.set_method(method)
        fit.initialize(x, y, srt=False)
        ys0 = fit.smooth(f, nsteps=nsteps)
        rw0, res0 = fit.m_weights.copy(), fit.m_residuals.copy()
        fit.close()
        cf.set_default_method(method)
        try:
            ys1, rw1, res1 = cf.cf_lowess_smooth(x, y, f, nsteps)
        finally:
            cf.set_default_method("direct")
        assert np.array_equal(ys1, ys0), (method, nsteps)
        assert np.array_equal(res1, res0), (method, nsteps)
        if nsteps > 0:
            assert np.array_equal(rw1, rw0), (method, nsteps)
