"""GPU parity tests of the moment-tree evaluator (CF_METHOD_TREE, tree.cu; SURVEY 8f
row N2) against the oracle, through the C ABI.  Same tolerance as the direct
kernel: 1e-10 (tests/cases.close).  The tree must also agree with the direct
kernel on sizes the oracle cannot run in full."""
import ctypes as C

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


def _tree_fit(cf, x, y):
    fit = cf.lowess_smoothing()
    fit.set_method("tree")
    fit.initialize(x, y, srt=False)
    return fit


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal", "wide"])
@pytest.mark.parametrize("n,f,nsteps", [(2, 0.5, 2), (3, 1.0, 0), (17, 1.0, 2), (100, 0.2, 2), (100, 0.8, 2),
                                        (1000, 0.05, 3), (4099, 0.3, 2), (513, 1.0, 1),
                                        (2048, 0.01, 2), (600, 0.4, 0), (3000, 0.2, 5),
                                        (16384, 0.3, 2), (20001, 0.9, 1)])
def test_tree_smooth_matches_oracle(cf, oracle, kind, n, f, nsteps):
    x = cases.make_x(kind, n, n + nsteps)
    y = cases.make_y(x, n + nsteps)
    ref = oracle.lowess_mt(x, y, f, nsteps) if n > 5000 else oracle.lowess(x, y, f, nsteps)
    fit = _tree_fit(cf, x, y)
    ys = fit.smooth(f, nsteps=nsteps)
    # pointwise 1e-10 against the oracle; the few points (zero crossings at n >= 16384)
    # where the oracle's OWN rounding exceeds that are refereed in extended precision
    ok, d, nref = cases.close_or_exact(ys, ref["ys"], x, y, fit.m_weights if nsteps > 0 else None, f, TOL)
    if x.size < 16384:
        assert nref == 0, f"{nref} point(s) needed the referee at n = {x.size}"
    assert ok, ("ys", d, nref)
    d = float(np.max(np.abs(fit.m_residuals - ref["res"])))
    assert d <= TOL * max(float(np.max(np.abs(y))), 1e-300), ("res", d)
    if nsteps > 0:
        assert np.max(np.abs(fit.m_weights - ref["rw"])) <= 1e-9, "rw"
        assert abs(fit.last_stats.cmad_last - ref["cmad"][-1]) <= 1e-10 * max(ref["cmad"][-1], 1e-300)


@pytest.mark.parametrize("yoff,yscale", [(1.0e3, 1.0), (0.0, 1.0e-6), (-50.0, 1.0e4)])
def test_tree_offset_and_scale_of_y(cf, oracle, yoff, yscale):
    n, f = 5000, 0.15
    x = cases.make_x("uniform", n, 7)
    y = cases.make_y(x, 7) * yscale + yoff
    ref = oracle.lowess(x, y, f, 2)
    ys = _tree_fit(cf, x, y).smooth(f)
    ok, d = cases.close(ys, ref["ys"], TOL)
    assert ok, d


def test_tree_zero_mad_and_dead_windows(cf, oracle):
    n = 600
    x = datagen.grid_x(n)
    y = np.zeros(n)
    y[200:206] = [5.0, -3.0, 7.0, -4.0, 6.0, -2.0]
    ref = oracle.lowess(x, y, 0.1, 2)
    assert ref["cmad"][-1] == 0.0
    fit = _tree_fit(cf, x, y)
    ys = fit.smooth(0.1)
    assert np.array_equal(fit.m_weights, ref["rw"])
    assert cases.close(ys, ref["ys"], TOL)[0]
    dead = np.nonzero(ref["ys"] == y)[0]
    assert np.array_equal(ys[dead], y[dead])


def test_tree_agrees_with_direct_kernel(cf):
    """n = 1e5, f = 0.05 and f = 0.5: both backends, full smooth, same tolerance."""
    n = 100000
    x = np.sort(datagen.random_x(n, 3))
    y = datagen.noisy_y(x, 3)
    for f in (0.05, 0.5):
        d = cf.lowess_smoothing()
        d.initialize(x, y, srt=False)
        ys_d = d.smooth(f, nsteps=2)
        t = _tree_fit(cf, x, y)
        ys_t = t.smooth(f, nsteps=2)
        ok, err = cases.close(ys_t, ys_d, TOL)
        assert ok, (f, err)
        assert np.max(np.abs(t.m_weights - d.m_weights)) <= 1e-9


def test_tree_batched(cf, oracle):
    s, n, f = 40, 1500, 0.3
    x = np.stack([np.sort(datagen.random_x(n, k)) for k in range(s)])
    y = np.stack([datagen.noisy_y(x[k], k) for k in range(s)])
    b = cf.batched_lowess()
    b.set_method("tree")
    b.initialize(x, y)
    b.run_resident(f, 2)
    ys, rw, res = b.fetch()
    for k in (0, 1, 17, 39):
        ref = oracle.lowess(x[k], y[k], f, 2)
        assert cases.close(ys[k], ref["ys"], TOL)[0], k
        assert np.max(np.abs(rw[k] - ref["rw"])) < 1e-9, k


def test_tree_shards_bitwise(cf):
    from curvefit_b200 import _lib
    n, f = 6000, 0.1
    x = np.sort(datagen.random_x(n, 2))
    y = datagen.noisy_y(x, 2)
    whole = _tree_fit(cf, x, y)
    ys_all, _, _ = whole.run_from_rw(f, None, 1)
    lib = _lib.load()
    for G in (2, 3):
        for g in range(G):
            part = _tree_fit(cf, x, y)
            assert lib.cf_lowess_set_shard(part._h, g, G) == 0
            b, e, blk = C.c_int32(), C.c_int32(), C.c_int32()
            lib.cf_lowess_shard(part._h, C.byref(b), C.byref(e), C.byref(blk))
            ys, _, _ = part.run_from_rw(f, None, 1)
            assert np.array_equal(ys[b.value:e.value], ys_all[b.value:e.value]), (G, g)


@pytest.mark.parametrize("n,f,nsteps", [(200000, 0.8, 5), (1000000, 0.05, 3), (3000000, 0.3, 2)])
def test_tree_large_teacher_forced(cf, oracle, n, f, nsteps):
    """Full-size runs the oracle cannot finish: its lowest() at sampled points, with the
    GPU's own robustness weights entering the pass (teacher forcing)."""
    x = datagen.grid_x(n)
    y = datagen.noisy_y(x)
    fit = _tree_fit(cf, x, y)
    ys = fit.smooth(f, nsteps=nsteps)
    assert np.all(np.isfinite(ys))
    rw_in = fit.m_weights.copy()
    ys_next, _, _ = fit.run_from_rw(f, rw_in, passes=1)
    m = 60 if n * f > 5e5 else 120
    idx = np.unique(np.concatenate([np.arange(0, 3), np.arange(n - 3, n),
                                    np.linspace(0, n - 1, m).astype(np.int64)])).astype(np.int32)
    t = oracle.lowest_at(x, y, f, idx, rw_in)
    ok, d, nref = cases.close_or_exact(ys_next[idx], t["ys"], x, y, rw_in, f, TOL, idx=idx)
    assert ok, (d, nref)
    assert np.all((rw_in >= 0.0) & (rw_in <= 1.0))
