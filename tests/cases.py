"""Seeded input families shared by the CPU and GPU parity tests."""
import numpy as np

from curvefit_b200 import datagen


def make_x(kind: str, n: int, seed: int = 0) -> np.ndarray:
    rng = np.random.default_rng(1234 + seed)
    if kind == "grid":
        return datagen.grid_x(n)
    if kind == "uniform":
        return np.sort(datagen.random_x(n, seed))
    if kind == "clustered":
        c = rng.normal(size=n) * 0.02 + rng.integers(0, 4, size=n) * 0.3
        return np.sort(c)
    if kind == "ties":
        return np.sort(np.round(rng.random(n) * max(n // 8, 2)) / max(n // 8, 2))
    if kind == "equal":
        return np.full(n, 0.75)
    if kind == "wide":
        return np.sort(np.exp(rng.normal(size=n) * 3.0))
    raise ValueError(kind)


def make_y(x: np.ndarray, seed: int = 0, outliers: bool = True) -> np.ndarray:
    y = datagen.noisy_y(x, seed)
    if outliers and x.size >= 20:
        rng = np.random.default_rng(99 + seed)
        k = rng.choice(x.size, size=max(x.size // 50, 1), replace=False)
        y[k] += rng.choice([-8.0, 8.0], size=k.size)
    return y


def close(a, b, tol=1e-10):
    """|d_i| <= tol * max(|ref_i|, 1e-3 * max|ref|) and normwise tol."""
    a = np.asarray(a)
    b = np.asarray(b)
    scale = np.max(np.abs(b)) if b.size else 0.0
    floor = np.maximum(np.abs(b), 1e-3 * scale)
    d = np.abs(a - b)
    point = bool(np.all(d <= tol * floor + 1e-300))
    norm = bool(d.max(initial=0.0) <= tol * scale + 1e-300)
    return point and norm, float(d.max(initial=0.0))


# every use of the extended-precision referee below: (test id, n, refereed points) -- printed in
# the pytest summary (tests/conftest.py) so that nobody has to guess how often it fires
REFEREE_LOG = []


def close_or_exact(got, ref, x, y, rw, f, tol=1e-10, max_referee=64, idx=None):
    """close(); where single points miss the pointwise bound, the extended-precision
    evaluation of the reference's own formula (py_lowess.lowest_extended, with the
    robustness weights `rw` that entered the pass; None = unweighted) is the referee:
    `got` must meet the same bound against IT.  The normwise bound against `ref` always
    applies.  `idx`: positions in x of the entries of got/ref (default: all points).
    Returns (ok, max |got - ref|, number of refereed points)."""
    from tests import py_lowess
    got = np.asarray(got)
    ref = np.asarray(ref)
    scale = np.max(np.abs(ref)) if ref.size else 0.0
    d = np.abs(got - ref)
    if not d.max(initial=0.0) <= tol * scale + 1e-300:
        return False, float(d.max(initial=0.0)), 0
    bad = np.nonzero(d > tol * np.maximum(np.abs(ref), 1e-3 * scale) + 1e-300)[0]
    if bad.size:
        import os
        REFEREE_LOG.append((os.environ.get("PYTEST_CURRENT_TEST", "?"), int(x.size), int(bad.size)))
    if bad.size > max_referee:
        return False, float(d.max(initial=0.0)), int(bad.size)
    n = x.size
    ns = max(min(int(f * float(n)), n), 2)
    for i in bad:
        ex = py_lowess.lowest_extended(x, y, rw, int(i if idx is None else idx[i]), ns)
        if not abs(got[i] - ex) <= tol * max(abs(ex), 1e-3 * scale):
            return False, float(abs(got[i] - ex)), int(bad.size)
    return True, float(d.max(initial=0.0)), int(bad.size)


# The worked example printed in the header of netlib's lowess.f (W. S. Cleveland, Bell
# Laboratories; "go/lowess.f"), the routine the reference's lowess / lowest are a Fortran 90
# port of (src/curvefit_regression.f90:508-512, :620-624 cite the same paper): twenty points,
# ten of them with equal abscissas, F = .25, and the smoothed values to three decimals for
# (NSTEPS, DELTA) = (0, 0), (0, 3), (2, 0).  A published known answer for the ALGORITHM --
# ties, the delta interpolation and the robustness iterations included; the reference itself
# holds no LOWESS fixture.
NETLIB_X = np.array([1, 2, 3, 4, 5] + [6] * 10 + [8, 10, 12, 14, 50], dtype=np.float64)
NETLIB_Y = np.array([18, 2, 15, 6, 10, 4, 16, 11, 7, 3, 14, 17, 20, 12, 9, 13, 1, 8, 5, 19], dtype=np.float64)
NETLIB_F = 0.25
NETLIB_YS = {
    (0, 0.0): [13.659, 11.145, 8.701, 9.722, 10.000] + [11.300] * 10 + [13.000, 6.440, 5.596, 5.456, 18.998],
    (0, 3.0): [13.659, 12.347, 11.034, 9.722, 10.511] + [11.300] * 10 + [13.000, 6.440, 5.596, 5.456, 18.998],
    (2, 0.0): [14.811, 12.115, 8.984, 9.676, 10.000] + [11.346] * 10 + [13.000, 6.734, 5.744, 5.415, 18.998],
}
