"""Numerical prototype of the moment-tree evaluator in numpy: same
node layout and formulas as curvefit_b200/csrc/tree.cu, used to check the
accuracy of the scheme against the oracle before/while writing the kernel.

    python scripts/proto_tree.py [n] [f]
"""
import sys
import os
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle import oracle          # noqa: E402  (development script; not product code)
from tests import cases            # noqa: E402

B = 16
KA, KB = 12, 11


def build(x, y, rw):
    n = x.size
    nleaf = (n + B - 1) // B
    levels = []
    # leaf level
    lo = np.arange(nleaf) * B
    hi = np.minimum(lo + B, n) - 1
    c = 0.5 * (x[lo] + x[hi])
    s = 0.5 * (x[hi] - x[lo])
    A = np.zeros((nleaf, KA))
    Bm = np.zeros((nleaf, KB))
    for k in range(nleaf):
        j = slice(lo[k], hi[k] + 1)
        inv = 1.0 / s[k] if s[k] > 0 else 0.0
        v = (x[j] - c[k]) * inv
        p = rw[j].copy()
        py = rw[j] * y[j]
        for m in range(KA):
            A[k, m] = p.sum()
            if m < KB:
                Bm[k, m] = py.sum()
            p = p * v
            py = py * v
    levels.append((c, s, A, Bm, lo, hi))
    from math import comb
    while levels[-1][0].size > 1:
        c0, s0, A0, B0, lo0, hi0 = levels[-1]
        m = (c0.size + 1) // 2
        lo1 = lo0[0::2]
        hi1 = np.array([hi0[min(2 * k + 1, c0.size - 1)] for k in range(m)])
        c1 = 0.5 * (x[lo1] + x[hi1])
        s1 = 0.5 * (x[hi1] - x[lo1])
        A1 = np.zeros((m, KA))
        B1 = np.zeros((m, KB))
        for k in range(m):
            inv = 1.0 / s1[k] if s1[k] > 0 else 0.0
            for ch in (2 * k, 2 * k + 1):
                if ch >= c0.size:
                    continue
                a = s0[ch] * inv
                b = (c0[ch] - c1[k]) * inv
                for kk in range(KA):
                    A1[k, kk] += sum(comb(kk, mm) * b ** (kk - mm) * a ** mm * A0[ch, mm] for mm in range(kk + 1))
                    if kk < KB:
                        B1[k, kk] += sum(comb(kk, mm) * b ** (kk - mm) * a ** mm * B0[ch, mm] for mm in range(kk + 1))
        levels.append((c1, s1, A1, B1, lo1, hi1))
    return levels


def take(levels, l, idx, xs, invh, sigma, core):
    c, s, A, Bm, _, _ = levels[l]
    beta = (c[idx] - xs) * invh
    alpha = s[idx] * invh
    A = A[idx]
    Bm = Bm[idx]
    if core:
        f = np.zeros(10)
        f[0] = 1.0
    else:
        e = np.array([beta ** 3, 3 * beta * beta * alpha, 3 * beta * alpha * alpha, alpha ** 3])
        cc = -sigma * e
        cc[0] += 1.0
        d = np.convolve(cc, cc)
        f = np.convolve(d, cc)
    D0 = f @ A[0:10]
    D1 = f @ A[1:11]
    D2 = f @ A[2:12]
    E0 = f @ Bm[0:10]
    E1 = f @ Bm[1:11]
    return np.array([D0, beta * D0 + alpha * D1, beta * beta * D0 + 2 * alpha * beta * D1 + alpha * alpha * D2,
                     E0, beta * E0 + alpha * E1])


def direct(x, y, rw, a, b, xs, invh, l1, r1):
    out = np.zeros(5)
    for j in range(a, b):
        q = (x[j] - xs) * invh
        w = (1.0 - q * q * abs(q)) ** 3
        if l1 <= j <= r1:
            w = 1.0
        wr = w * rw[j]
        out += np.array([wr, wr * q, wr * q * q, wr * y[j], wr * q * y[j]])
    return out


def rng_nodes(a, b):
    """[a,b) -> head direct [a,h), node leaves, tail direct [t,b)."""
    up = (a + B - 1) // B * B
    dn = b // B * B
    h = min(b, up)
    t = max(h, dn)
    nodes = []
    la, lb = up // B, dn // B
    l = 0
    while la < lb:
        if la & 1:
            nodes.append((l, la))
            la += 1
        if lb & 1:
            lb -= 1
            nodes.append((l, lb))
        la >>= 1
        lb >>= 1
        l += 1
    return h, nodes, t


def evaluate(x, y, rw, levels, i, ns):
    n = x.size
    xs = x[i]
    # windows as windows.cu
    nl = 0
    while nl < n - ns and not (xs - x[nl] <= x[nl + ns] - xs):
        nl += 1
    nr = nl + ns - 1
    h = max(xs - x[nl], x[nr] - xs)
    h9, h1 = 0.999 * h, 0.001 * h
    r = np.abs(x - xs)
    idx = np.arange(n)
    in9 = (r <= h9) & (idx >= nl)
    # contiguous from nl side: first sweep stops at first x > xs with r > h9
    l9 = int(np.argmax(in9))
    r9 = i
    while r9 + 1 < n and r[r9 + 1] <= h9:
        r9 += 1
    l1 = i
    while l1 - 1 >= l9 and r[l1 - 1] <= h1:
        l1 -= 1
    r1 = i
    while r1 + 1 <= r9 and r[r1 + 1] <= h1:
        r1 += 1
    invh = 1.0 / h if h > 0 else 0.0
    S = np.zeros(5)
    cnt = 0
    for (a, b, sigma, core) in ((l9, l1, -1.0, False), (l1, r1 + 1, 0.0, True), (r1 + 1, r9 + 1, 1.0, False)):
        hd, nodes, tl = rng_nodes(a, b)
        S += direct(x, y, rw, a, hd, xs, invh, l1, r1)
        S += direct(x, y, rw, tl, b, xs, invh, l1, r1)
        for (l, k) in nodes:
            S += take(levels, l, k, xs, invh, sigma, core)
            cnt += 1
    S0, S1, S2, T0, T1 = S
    if not S0 > 0:
        return y[i], cnt
    ym = T0 / S0
    out = ym
    if h > 0:
        um = S1 / S0
        cu = S2 / S0 - um * um
        cu = max(cu, 0.0)
        if np.sqrt(cu) * h > 0.001 * (x[-1] - x[0]):
            cov = T1 / S0 - um * ym
            out = ym + cov / cu * (0.0 - um)
    return out, cnt


def main():
    n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 4000
    f = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
    rng = np.random.default_rng(5)
    for kind in ("grid", "uniform", "clustered", "ties", "wide", "equal"):
        x = cases.make_x(kind, n)
        y = cases.make_y(x)
        for yoff in (0.0, 1e3):
            yy = y + yoff
            rw = rng.random(n)
            rw[rng.random(n) < 0.1] = 0.0
            levels = build(x, yy, rw)
            ns = oracle.span(n, f)
            pts = np.unique(np.concatenate([np.arange(0, 5), np.arange(n - 5, n), rng.integers(0, n, 150)])).astype(np.int32)
            ref = oracle.lowest_at(x, yy, f, pts, rw)["ys"]
            got = np.array([evaluate(x, yy, rw, levels, int(i), ns)[0] for i in pts])
            nodes = np.mean([evaluate(x, yy, rw, levels, int(i), ns)[1] for i in pts[:20]])
            err = np.max(np.abs(got - ref))
            scale = np.max(np.abs(ref))
            print(f"{kind:10s} yoff={yoff:6.0f} n={n} ns={ns} max|d|={err:.3e} rel={err / scale:.3e} nodes/pt={nodes:.1f}")


if __name__ == "__main__":
    main()
