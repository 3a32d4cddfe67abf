"""Independent pure-Python transliteration of lowess/lowest
(src/curvefit_regression.f90:513-586, :625-715), written from the Fortran with
1-based lists and Python floats (IEEE double, no FMA).  Test infrastructure:
it exists only to cross-check oracle/lowess_oracle.c on small inputs.
"""
import math
import sys


def lowest(x, y, n, xs, nleft, nright, w, userw, rw):
    """Returns (ys, ok, nrt); x, y, w, rw are 1-based lists (index 0 unused)."""
    rng = x[n] - x[1]
    h = max(xs - x[nleft], x[nright] - xs)
    h9 = 0.999 * h
    h1 = 0.001 * h
    a = 0.0
    j = nleft
    while j <= n:
        w[j] = 0.0
        r = abs(x[j] - xs)
        if r <= h9:
            if r > h1:
                q = r / h
                q3 = q * q * q
                t = 1.0 - q3
                w[j] = t * t * t
            else:
                w[j] = 1.0
            if userw:
                w[j] = rw[j] * w[j]
            a = a + w[j]
        elif x[j] > xs:
            break
        j += 1
    nrt = j - 1
    if a <= 0.0:
        return None, False, nrt
    for j in range(nleft, nrt + 1):
        w[j] = w[j] / a
    if h > 0.0:
        a = 0.0
        for j in range(nleft, nrt + 1):
            a = a + w[j] * x[j]
        b = xs - a
        c = 0.0
        for j in range(nleft, nrt + 1):
            d = x[j] - a
            c = c + w[j] * (d * d)
        if math.sqrt(c) > 0.001 * rng:
            b = b / c
            for j in range(nleft, nrt + 1):
                w[j] = w[j] * (1.0 + b * (x[j] - a))
    ys = 0.0
    for j in range(nleft, nrt + 1):
        ys = ys + w[j] * y[j]
    return ys, True, nrt


def lowess(x0, y0, f, nsteps, delta=0.0):
    """Returns (ys, rw, res, windows) as 0-based lists; windows[i] = (nleft, nright, nrt)
    (1-based values) for fitted points, None otherwise."""
    n = len(x0)
    x = [0.0] + [float(v) for v in x0]
    y = [0.0] + [float(v) for v in y0]
    ys = [0.0] * (n + 1)
    rw = [0.0] * (n + 1)
    res = [0.0] * (n + 1)
    win = [None] * (n + 1)
    ns = max(min(int(f * float(n)), n), 2)
    eps = sys.float_info.epsilon
    if n < 2:
        return y[1:], rw[1:], res[1:], win[1:]
    for it in range(1, nsteps + 2):
        nleft, nright, last, i = 1, ns, 0, 1
        while True:
            while nright < n:
                d1 = x[i] - x[nleft]
                d2 = x[nright + 1] - x[i]
                if d1 <= d2:
                    break
                nleft += 1
                nright += 1
            v, ok, nrt = lowest(x, y, n, x[i], nleft, nright, res, it > 1, rw)
            win[i] = (nleft, nright, nrt)
            ys[i] = v if ok else y[i]
            if last < i - 1:
                denom = x[i] - x[last]
                for j in range(last + 1, i):
                    alpha = (x[j] - x[last]) / denom
                    ys[j] = alpha * ys[i] + (1.0 - alpha) * ys[last]
            last = i
            cut = x[last] + delta
            i = last + 1
            while i <= n:
                if x[i] > cut:
                    break
                if abs(x[i] - x[last]) < eps:
                    ys[i] = ys[last]
                    last = i
                i += 1
            i = max(last + 1, i - 1)
            if last >= n:
                break
        for k in range(1, n + 1):
            res[k] = y[k] - ys[k]
        if it > nsteps:
            break
        srt = sorted(abs(v) for v in res[1:])
        srt = [0.0] + srt
        m1 = 1 + n // 2
        m2 = n - m1 + 1
        cmad = 3.0 * (srt[m1] + srt[m2])
        c9 = 0.999 * cmad
        c1 = 0.001 * cmad
        for k in range(1, n + 1):
            r = abs(res[k])
            if r <= c1:
                rw[k] = 1.0
            elif r > c9:
                rw[k] = 0.0
            else:
                q = r / cmad
                t = 1.0 - q * q
                rw[k] = t * t
    return ys[1:], rw[1:], res[1:], win[1:]


def lowest_extended(x, y, rw, i, ns):
    """The value lowest() (:513-586) would return at 0-based point i in (nearly) EXACT
    arithmetic: every discrete decision (window slide, r <= h9, r <= h1) is taken in
    IEEE double exactly as the reference takes it, the sums are then formed in x87
    extended precision (64-bit mantissa) with numpy.  Referee for points where two
    double-precision evaluations differ by more than the tolerance because of their
    own rounding (the reference's sequential sums carry ~1e-13 at ns ~ 5000)."""
    import numpy as np
    L = np.longdouble
    n = x.size
    xs = x[i]
    cond = (xs - x[:n - ns]) <= (x[ns:] - xs)      # slide predicate d1 <= d2, :664-666
    nl = int(np.argmax(cond)) if cond.any() else n - ns
    nr = nl + ns - 1
    h = max(xs - x[nl], x[nr] - xs)
    h9, h1 = 0.999 * h, 0.001 * h
    r = np.abs(x - xs)
    out = np.nonzero((r[i:] > h9) & (x[i:] > xs))[0]   # sweep exit: first x_j > xs with r > h9
    stop = i + int(out[0]) if out.size else n
    m = np.zeros(n, bool)
    m[nl:stop] = r[nl:stop] <= h9
    xm = x[m].astype(L)
    w = np.ones(xm.size, L)
    if h > 0.0:
        q = np.abs(xm - L(xs)) / L(h)
        w = (1 - q ** 3) ** 3
        w[r[m] <= h1] = 1
    if rw is not None:
        w = w * rw[m].astype(L)
    a = w.sum()
    if not a > 0:
        return float(y[i])
    w = w / a
    if h > 0.0:
        mean = (w * xm).sum()
        b = L(xs) - mean
        c = (w * (xm - mean) ** 2).sum()
        if np.sqrt(c) > L(0.001) * (L(x[-1]) - L(x[0])):
            w = w * (1 + (b / c) * (xm - mean))
    return float((w * y[m].astype(L)).sum())
