"""CPU tests of the oracle (oracle/lowess_oracle.c): the restatement is checked
against an independent pure-Python transliteration, analytic properties, a
dense weighted-least-squares solve, and our frozen golden vectors.  The
reference's own LOWESS test asserts nothing
(/root/reference tests/curvefit_test_regression.f90:10-48), so these stand in
for goldens (PARITY UNPINNED, see oracle header)."""
import os
import sys

import numpy as np
import pytest

from curvefit_b200 import datagen
from tests import cases, py_lowess

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal", "wide"])
@pytest.mark.parametrize("n,f,nsteps", [(2, 0.5, 2), (3, 1.0, 1), (37, 0.2, 2), (100, 0.8, 2),
                                        (150, 0.05, 3), (64, 0.01, 0)])
def test_c_oracle_matches_python_transliteration_bitwise(oracle, kind, n, f, nsteps):
    x = cases.make_x(kind, n, seed=n)
    y = cases.make_y(x, seed=n)
    r = oracle.lowess(x, y, f, nsteps, want_windows=True)
    ys, rw, res, win = py_lowess.lowess(x, y, f, nsteps)
    assert np.array_equal(r["ys"], np.array(ys))
    assert np.array_equal(r["res"], np.array(res))
    if nsteps > 0:
        assert np.array_equal(r["rw"], np.array(rw))
    for i, w in enumerate(win):
        if w is None:
            assert r["nleft"][i] == 0
        else:
            assert (r["nleft"][i], r["nright"][i], r["nrt"][i]) == w


@pytest.mark.parametrize("delta", [0.01, 0.05])
def test_delta_interpolation_matches_python(oracle, delta):
    x = cases.make_x("uniform", 200, 5)
    y = cases.make_y(x, 5)
    r = oracle.lowess(x, y, 0.3, 2, delta=delta)
    ys, rw, res, _ = py_lowess.lowess(x, y, 0.3, 2, delta)
    assert np.array_equal(r["ys"], np.array(ys))
    assert np.array_equal(r["rw"], np.array(rw))


@pytest.mark.parametrize("kind", ["grid", "uniform", "clustered", "ties", "equal"])
@pytest.mark.parametrize("n,f", [(100, 0.2), (1000, 0.05), (1000, 1.0), (513, 0.3), (50, 0.0)])
def test_binary_search_window_equals_sequential_slide(oracle, kind, n, f):
    """F9: the slide loop (:663-669) equals an independent search per point."""
    x = cases.make_x(kind, n, 11)
    y = cases.make_y(x, 11)
    r = oracle.lowess(x, y, f, 0, want_windows=True)
    fitted = np.nonzero(r["nleft"])[0].astype(np.int32)
    t = oracle.lowest_at(x, y, f, fitted)
    assert np.array_equal(t["nleft"], r["nleft"][fitted])
    assert np.array_equal(t["nright"], r["nright"][fitted])
    assert np.array_equal(t["nrt"], r["nrt"][fitted])
    assert np.array_equal(t["ys"], r["ys"][fitted])


@pytest.mark.parametrize("kind", ["grid", "uniform", "ties", "equal"])
def test_multithreaded_oracle_is_bit_identical(oracle, kind):
    x = cases.make_x(kind, 700, 2)
    y = cases.make_y(x, 2)
    a = oracle.lowess(x, y, 0.15, 3)
    b = oracle.lowess_mt(x, y, 0.15, 3, nthreads=4)
    for k in ("ys", "rw", "res", "cmad"):
        assert np.array_equal(a[k], b[k]), k


def test_exactly_linear_data_is_reproduced(oracle):
    x = cases.make_x("uniform", 400, 1)
    y = 3.0 - 2.0 * x
    r = oracle.lowess(x, y, 0.25, 2)
    assert np.max(np.abs(r["ys"] - y)) < 1e-12
    # residuals are rounding noise, so cmad ~ 1e-16 and the weights are arbitrary;
    # the fit must stay on the line regardless
    r0 = oracle.lowess(x, y, 0.25, 0)
    assert np.max(np.abs(r0["ys"] - y)) < 1e-12


def test_constant_y(oracle):
    x = cases.make_x("clustered", 300, 4)
    y = np.full(300, 2.5)
    r = oracle.lowess(x, y, 0.4, 2)
    assert np.max(np.abs(r["ys"] - 2.5)) < 1e-13


def test_first_pass_equals_dense_weighted_least_squares(oracle):
    """Non-robust pass vs an independent numpy WLS solve per point."""
    n, f = 200, 0.3
    x = cases.make_x("uniform", n, 9)
    y = cases.make_y(x, 9, outliers=False)
    r = oracle.lowess(x, y, f, 0, want_windows=True)
    ns = oracle.span(n, f)
    for i in range(n):
        nl, nr = r["nleft"][i] - 1, r["nright"][i] - 1
        h = max(x[i] - x[nl], x[nr] - x[i])
        d = np.abs(x - x[i])
        w = np.where(d <= 0.001 * h, 1.0, (1 - (d / h) ** 3) ** 3)
        w[d > 0.999 * h] = 0.0
        w[:nl] = 0.0
        A = np.stack([np.ones(n), x - x[i]], axis=1)
        sw = np.sqrt(w)
        coef, *_ = np.linalg.lstsq(A * sw[:, None], y * sw, rcond=None)
        assert abs(coef[0] - r["ys"][i]) < 1e-9, i
    assert ns == 60


def test_span_is_truncation(oracle):
    # F3: int(f*n), not ceil; values verified in SURVEY section 0
    assert oracle.span(16384, 0.3) == 4915
    assert oracle.span(100000000, 0.05) == 5000000
    assert oracle.span(1000000, 0.05) == 50000
    assert oracle.span(2000000, 0.8) == 1600000
    assert oracle.span(100, 0.2) == 20 and oracle.span(100, 0.8) == 80
    assert oracle.span(100, 0.0) == 2 and oracle.span(100, 5.0) == 100


def test_quick_return_n_lt_2(oracle):
    r = oracle.lowess(np.array([1.0]), np.array([7.0]), 0.5, 2)
    assert r["ys"][0] == 7.0 and r["rw"][0] == 0.0 and r["res"][0] == 0.0


def test_zero_mad_and_all_zero_weight_windows(oracle):
    # more than half the residuals exactly zero -> cmad = 0 -> rw in {0, 1} (:706-711);
    # windows whose weights are all zero -> ok = .false. -> ys = y (:559, :673)
    n = 60
    x = datagen.grid_x(n)
    y = np.zeros(n)
    y[20:26] = 5.0
    r = oracle.lowess(x, y, 0.1, 2)
    assert set(np.unique(r["rw"])) <= {0.0, 1.0}
    assert np.all(np.isfinite(r["ys"]))
    p = py_lowess.lowess(x, y, 0.1, 2)
    assert np.array_equal(r["ys"], np.array(p[0]))


def test_index_sort_matches_reference_algorithm(oracle):
    """LINALG's own test (tests/test_sort.f90:33-54) checks values only; so do we,
    plus that ind is the permutation that produced them."""
    rng = np.random.default_rng(5)
    for n in (1, 2, 200, 5000):
        x = rng.random(n)
        xs, ind = oracle.sort_ind(x)
        assert np.array_equal(xs, np.sort(x))
        assert np.array_equal(x[ind - 1], xs)
        assert np.array_equal(np.sort(ind), np.arange(1, n + 1))
    x = np.round(rng.random(500) * 10) / 10          # ties: unstable order allowed
    xs, ind = oracle.sort_ind(x)
    assert np.array_equal(xs, np.sort(x)) and np.array_equal(x[ind - 1], xs)
    xs, ind = oracle.sort_ind(np.arange(3000.0))      # sorted input: deep recursion path
    assert np.array_equal(xs, np.arange(3000.0))


def test_ls_init(oracle):
    rng = np.random.default_rng(6)
    x = rng.random(300)
    y = rng.random(300)
    mx, my = oracle.ls_init(x, y, True)
    o = np.argsort(x, kind="stable")
    assert np.array_equal(mx, x[o]) and np.array_equal(my, y[o])
    mx, my = oracle.ls_init(x, y, False)
    assert np.array_equal(mx, x) and np.array_equal(my, y)


@pytest.mark.parametrize("name", ["c1_example2", "n4096_uniform", "n1000_ties"])
def test_oracle_reproduces_golden_vectors(oracle, name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    if name == "c1_example2":
        x, y = datagen.example2(100)
        assert np.array_equal(x, g["x"]) and np.array_equal(y, g["y"])
        for f, tag in ((0.2, "f02"), (0.8, "f08")):
            r = oracle.lowess(x, y, f, 2, want_windows=True)
            for k in ("ys", "rw", "res", "cmad", "nleft", "nright", "nrt"):
                assert np.array_equal(r[k], g[f"{tag}_{k}"]), (tag, k)
    else:
        r = oracle.lowess(g["x"], g["y"], float(g["f"]), int(g["nsteps"]), want_windows=True)
        for k in ("ys", "rw", "res", "cmad", "nleft", "nright", "nrt"):
            assert np.array_equal(r[k], g[k]), k


# ---- neighbours of the path: moving_average, median (oracle/stats_oracle.c) ----------------
def test_median_oracle_known_answer(oracle):
    """The reference's own known-answer test, tests/curvefit_test_statistics.f90:286-309."""
    z, xs = oracle.median([2.0, 3.0, 1.0, -4.0, -7.0, 1.0, 3.0])
    assert z == 1.0
    assert list(xs) == [-7.0, -4.0, 1.0, 1.0, 2.0, 3.0, 3.0]
    assert oracle.median([])[0] == 0.0
    assert oracle.median([4.0, 1.0])[0] == 2.5
    assert oracle.median([1.0, 4.0, 9.0], srt=False)[0] == 4.0


def test_moving_average_oracle_against_python_transliteration(oracle):
    """Independent restatement of src/curvefit_regression.f90:467-477 with mean_dbl
    (src/curvefit_statistics.f90:239-242) in plain Python floats (IEEE double)."""
    rng = np.random.default_rng(8)
    for n, npts in [(2, 2), (12, 3), (50, 50), (200, 17)]:
        x = rng.normal(size=n)
        buf = [0.0] * npts
        want = []
        for v in x:
            buf = [float(v)] + buf[:-1]
            z = buf[0]
            for k in range(2, npts + 1):
                z = z + (buf[k - 1] - z) / k
            want.append(z)
        got, rc = oracle.moving_average(x, npts)
        assert rc == 0 and np.array_equal(got, np.array(want))
    assert oracle.moving_average(np.zeros(5), 1)[1] == 201
    assert oracle.moving_average(np.zeros(5), 6)[1] == 201


@pytest.mark.parametrize("nsteps,delta", sorted(cases.NETLIB_YS))
def test_oracle_reproduces_the_published_netlib_example(oracle, nsteps, delta):
    """Known-answer test of the algorithm the reference ports (tests/cases.NETLIB_*): the
    oracle must print the same three decimals as the header of netlib's lowess.f."""
    r = oracle.lowess(cases.NETLIB_X, cases.NETLIB_Y, cases.NETLIB_F, nsteps, delta)
    want = np.array(cases.NETLIB_YS[(nsteps, delta)])
    assert np.array_equal(np.round(r["ys"], 3), want), np.round(r["ys"], 3) - want
    assert np.array_equal(r["res"], cases.NETLIB_Y - r["ys"])


@pytest.mark.parametrize("nsteps,delta", sorted(cases.NETLIB_YS))
def test_python_transliteration_reproduces_the_published_netlib_example(nsteps, delta):
    ys = py_lowess.lowess(cases.NETLIB_X, cases.NETLIB_Y, cases.NETLIB_F, nsteps, delta)[0]
    assert np.array_equal(np.round(np.asarray(ys, dtype=np.float64), 3), np.array(cases.NETLIB_YS[(nsteps, delta)]))


# ---- the untouched reference, wherever it can be built (oracle/ref/build_ref.sh) -----------------
def _ref_dump(name):
    import struct
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref",
                        f"out_{name}.bin")
    if not os.path.exists(path):
        return None
    raw = open(path, "rb").read()
    (n,) = struct.unpack("<i", raw[:4])
    a = np.frombuffer(raw, dtype="<f8", offset=4, count=4 * n).reshape(4, n)
    return {"xs": a[0], "ysrt": a[1], "ys": a[2], "res": a[3]}


@pytest.mark.parametrize("name", ["c1_f02", "c1_f08", "n4096_uniform", "n1000_ties", "n1500_clustered_srt"])
def test_oracle_matches_reference_dumps(oracle, name):
    """Pins the oracle on the reference ITSELF: smooth() of the untouched Fortran, built by
    oracle/ref/build_ref.sh and run on the inputs of tests/golden/make_ref_inputs.py.  Skips
    where no Fortran compiler has produced the dumps (this image, the B200 boxes): until it
    has run somewhere, parity stays 'unpinned' (DESIGN.md section 5)."""
    ref = _ref_dump(name)
    if ref is None:
        pytest.skip("oracle/_ref has no dump of the reference (no Fortran compiler here)")
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "golden"))
    import make_ref_inputs
    x, y, f, srt = make_ref_inputs.make(name)
    if srt:
        xs, ysrt = oracle.ls_init(x, y)
        assert np.array_equal(xs, ref["xs"])
        assert np.array_equal(ysrt, ref["ysrt"])          # same quicksort, same tie permutation
    else:
        xs, ysrt = x, y
    got = oracle.lowess(xs, ysrt, f, 2)
    assert np.array_equal(got["ys"], ref["ys"])
    assert np.array_equal(got["res"], ref["res"])
