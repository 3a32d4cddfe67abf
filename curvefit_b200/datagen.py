"""Deterministic synthetic LOWESS inputs.

The reference's only LOWESS data recipe is its (non-asserting) test,
tests/curvefit_test_regression.f90:26-33: x by cumulative addition of dx,
y = 0.5 sin(20x) + cos(5x) exp(-0.1x) + (u - 0.5) with unseeded u ~ U[0,1).
Here u comes from splitmix64 so that every run sees identical bits.
"""
from __future__ import annotations

import numpy as np

SEED_Y = 0x5EEDC0DE0000
SEED_X = 0x5EEDC0DE1000
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(z: np.ndarray) -> np.ndarray:
    """One splitmix64 output per uint64 counter value (vectorised)."""
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def uniform01(seed: int, n: int, start: int = 0) -> np.ndarray:
    """u_i = (splitmix64(seed + i) >> 11) * 2^-53, i = start .. start+n-1."""
    with np.errstate(over="ignore"):
        ctr = np.uint64(seed) + np.arange(start, start + n, dtype=np.uint64)
    return (splitmix64(ctr) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def signal(x: np.ndarray) -> np.ndarray:
    """tests/curvefit_test_regression.f90:31 / README.md:92."""
    return 0.5 * np.sin(20.0 * x) + np.cos(5.0 * x) * np.exp(-0.1 * x)


def grid_x(n: int, cumulative: bool | None = None) -> np.ndarray:
    """x on [0, 1]: cumulative x(i) = x(i-1) + dx (the reference's recipe,
    tests/curvefit_test_regression.f90:26-30) for small n, (i-1)*dx otherwise."""
    if n < 2:
        return np.zeros(n)
    dx = 1.0 / (n - 1.0)
    if cumulative is None:
        cumulative = n <= 100000
    if cumulative:
        x = np.empty(n)
        acc = 0.0
        for i in range(n):
            x[i] = acc
            acc = acc + dx
        return x
    return np.arange(n, dtype=np.float64) * dx


def random_x(n: int, series_id: int = 0) -> np.ndarray:
    """Unsorted, distinct-with-probability-1 abscissae in [0, 1)."""
    return uniform01(SEED_X + series_id, n)


def noisy_y(x: np.ndarray, series_id: int = 0) -> np.ndarray:
    """y = signal(x) + (u - 0.5), tests/curvefit_test_regression.f90:31-33."""
    return signal(x) + (uniform01(SEED_Y + series_id, x.size) - 0.5)


def example2(n: int = 100):
    """README Example 2 / test_lowess data with seeded noise."""
    x = grid_x(n, cumulative=True)
    return x, noisy_y(x)
