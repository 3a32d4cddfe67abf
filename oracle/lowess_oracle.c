/*
 * oracle/lowess_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (plain C) of the LOWESS hot path of jchristopherson/curvefit.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this.  The product (libcurvefit_cuda) never links,
 * loads or calls it.
 *
 * PARITY UNPINNED: the reference holds no golden vector or asserting test for
 * this path (tests/curvefit_test_regression.f90:10-48 asserts nothing and uses
 * unseeded noise) and cannot be compiled in this image (no Fortran compiler of
 * any kind).  This file is therefore a line-by-line restatement of the
 * reference's arithmetic, cross-checked against an independent pure-Python
 * transliteration (tests/py_lowess.py) and analytic properties.  The one
 * EXTERNAL known answer it is held to is the worked example in the header of
 * netlib's lowess.f -- the routine the reference's lowess/lowest port
 * (src/curvefit_regression.f90:508-512, :620-624 cite the same paper): twenty
 * points, ten tied, F = .25, (NSTEPS, DELTA) = (0, 0), (0, 3), (2, 0), reproduced
 * to the three printed decimals (tests/cases.NETLIB_*, tests/test_oracle.py).
 * That anchors the algorithm, not the reference's build of it: still unpinned.
 *
 * Every function cites the reference lines it follows.  Paths are relative to
 * /root/reference; LINALG = src/external/nonlin/src/external/linalg.
 *
 * Arithmetic contract (reference CMakeLists.txt:114: -O3 -funroll-all-loops,
 * no fast-math, baseline x86-64 => strict IEEE double, no FMA contraction):
 * build this file with -ffp-contract=off and without -ffast-math.
 * Integer powers follow gfortran's expansion: r**3 = (r*r)*r, r**2 = r*r.
 *
 * All arrays at this C boundary are 0-based; inside, pointers are shifted by
 * one so that the code reads with the reference's 1-based indices.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

/* error codes: src/curvefit_core.f90:34-40 */
#define CF_INVALID_INPUT_ERROR 201
#define CF_ARRAY_SIZE_ERROR 202
#define CF_OUT_OF_MEMORY_ERROR 105
#define CF_NO_DATA_DEFINED_ERROR 300

/* ---------------------------------------------------------------------------
 * lowest: src/curvefit_regression.f90:513-586.
 * x, y, rw, w are 1-based (already shifted).  Returns ok; *nrt_out gets the
 * last index the first sweep visited (:558).
 * ------------------------------------------------------------------------- */
static int lowest_1b(const double *x, const double *y, int n, double xs,
                     double *ys, int nleft, int nright, double *w, int userw,
                     const double *rw, int *nrt_out)
{
    const double zero = 0.0, one = 1.0, p001 = 1.0e-3, p999 = 0.999;
    int j, nrt;
    double range, h, h9, h1, a, b, c, r, d1, d2, t;

    range = x[n] - x[1];                                   /* :535 */
    d1 = xs - x[nleft];
    d2 = x[nright] - xs;
    h = d1 > d2 ? d1 : d2;                                 /* :536 max() */
    h9 = p999 * h;                                         /* :537 */
    h1 = p001 * h;                                         /* :538 */
    a = zero;

    for (j = nleft; j <= n; ++j) {                         /* :542 */
        w[j] = zero;
        r = fabs(x[j] - xs);
        if (r <= h9) {
            if (r > h1) {
                t = r / h;
                t = one - (t * t) * t;                     /* (r/h)**3 */
                w[j] = (t * t) * t;                        /* (...)**3 */
            } else {
                w[j] = one;
            }
            if (userw) w[j] = rw[j] * w[j];                /* :551 */
            a = a + w[j];
        } else if (x[j] > xs) {
            break;                                         /* :554 */
        }
    }
    nrt = j - 1;                                           /* :558 */
    if (nrt_out) *nrt_out = nrt;
    if (a <= zero) return 0;                               /* :559 */

    for (j = nleft; j <= nrt; ++j) w[j] = w[j] / a;        /* :563 */
    if (h > zero) {                                        /* :564 */
        a = zero;
        for (j = nleft; j <= nrt; ++j) a = a + w[j] * x[j];
        b = xs - a;
        c = zero;
        for (j = nleft; j <= nrt; ++j) {
            t = x[j] - a;
            c = c + w[j] * (t * t);                        /* :571 */
        }
        if (sqrt(c) > p001 * range) {                      /* :573 */
            b = b / c;
            for (j = nleft; j <= nrt; ++j)
                w[j] = w[j] * (one + b * (x[j] - a));      /* :576 */
        }
    }
    t = zero;
    for (j = nleft; j <= nrt; ++j) t = t + w[j] * y[j];    /* :581-584 */
    *ys = t;
    return 1;
}

static int cmp_dbl(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

/* span: src/curvefit_regression.f90:647  ns = max(min(int(f*n), n), 2) */
ORACLE_API int cf_oracle_span(int n, double f)
{
    double fn = f * (double)n;
    int k;
    if (!(fn < 2147483647.0)) k = n;        /* clamp before the int() cast */
    else if (fn < -2147483647.0) k = 2;
    else k = (int)fn;                       /* truncation toward zero */
    if (k > n) k = n;
    if (k < 2) k = 2;
    return k;
}

/* robustness weights: src/curvefit_regression.f90:697-713.
 * res, rw 1-based, n elements.  The full ascending sort (LINALG/src/
 * linalg_sorting.f90:12-35 -> LAPACK/SRC/dlasrt.f) is replaced by libc qsort:
 * sorted VALUES are algorithm independent.  Returns cmad. */
static double robust_weights_1b(const double *res, double *rw, int n)
{
    const double zero = 0.0, one = 1.0, three = 3.0, p001 = 1.0e-3, p999 = 0.999;
    int i, m1, m2;
    double cmad, c9, c1, r, t;
    for (i = 1; i <= n; ++i) rw[i] = fabs(res[i]);         /* :697 */
    qsort(rw + 1, (size_t)n, sizeof(double), cmp_dbl);     /* :698 */
    m1 = 1 + n / 2;                                        /* :699 */
    m2 = n - m1 + 1;                                       /* :700 */
    cmad = three * (rw[m1] + rw[m2]);                      /* :701 */
    c9 = p999 * cmad;
    c1 = p001 * cmad;
    for (i = 1; i <= n; ++i) {                             /* :704-713 */
        r = fabs(res[i]);
        if (r <= c1) rw[i] = one;
        else if (r > c9) rw[i] = zero;
        else {
            t = r / cmad;
            t = one - t * t;
            rw[i] = t * t;
        }
    }
    return cmad;
}

/* ---------------------------------------------------------------------------
 * lowess: src/curvefit_regression.f90:625-715 (sequential, the reference's
 * execution model).  Optional int outputs (may be NULL, length n, 0-based,
 * holding 1-based reference indices): nleft/nright/nrt of every point that
 * lowest() was called on in the LAST pass, 0 for tie-copied / interpolated
 * points.  cmad_out (length nsteps, may be NULL) gets each iteration's cmad.
 * ------------------------------------------------------------------------- */
ORACLE_API int cf_oracle_lowess(const double *x0, const double *y0, int n,
                                double f, int nsteps, double delta,
                                double *ys0, double *rw0, double *res0,
                                int *nleft_out, int *nright_out, int *nrt_out,
                                double *cmad_out)
{
    const double one = 1.0;
    const double *x = x0 - 1, *y = y0 - 1;
    double *ys = ys0 - 1, *rw = rw0 - 1, *res = res0 - 1;
    int ok, iter, i, j, nleft, nright, ns, last, nrt;
    double d1, d2, denom, alpha, cut, eps;

    ns = cf_oracle_span(n, f);                             /* :647 */
    eps = DBL_EPSILON;                                     /* :648 */
    if (n < 2) {                                           /* :651-654 */
        for (i = 1; i <= n; ++i) ys[i] = y[i];
        return 0;
    }
    if (nleft_out) memset(nleft_out, 0, sizeof(int) * (size_t)n);
    if (nright_out) memset(nright_out, 0, sizeof(int) * (size_t)n);
    if (nrt_out) memset(nrt_out, 0, sizeof(int) * (size_t)n);

    for (iter = 1; iter <= nsteps + 1; ++iter) {           /* :657 */
        nleft = 1;
        nright = ns;
        last = 0;
        i = 1;
        for (;;) {                                         /* :662 */
            while (nright < n) {                           /* :663-669 */
                d1 = x[i] - x[nleft];
                d2 = x[nright + 1] - x[i];
                if (d1 <= d2) break;
                nleft = nleft + 1;
                nright = nright + 1;
            }
            ok = lowest_1b(x, y, n, x[i], &ys[i], nleft, nright, res,
                           iter > 1, rw, &nrt);            /* :671 */
            if (nleft_out) nleft_out[i - 1] = nleft;
            if (nright_out) nright_out[i - 1] = nright;
            if (nrt_out) nrt_out[i - 1] = nrt;
            if (!ok) ys[i] = y[i];                         /* :673 */
            if (last < i - 1) {                            /* :674-680 */
                denom = x[i] - x[last];
                for (j = last + 1; j <= i - 1; ++j) {
                    alpha = (x[j] - x[last]) / denom;
                    ys[j] = alpha * ys[i] + (one - alpha) * ys[last];
                }
            }
            last = i;                                      /* :681 */
            cut = x[last] + delta;
            for (i = last + 1; i <= n; ++i) {              /* :683-689 */
                if (x[i] > cut) break;
                if (fabs(x[i] - x[last]) < eps) {
                    ys[i] = ys[last];
                    last = i;
                }
            }
            i = (last + 1 > i - 1) ? last + 1 : i - 1;     /* :690 */
            if (last >= n) break;                          /* :692 */
        }
        for (i = 1; i <= n; ++i) res[i] = y[i] - ys[i];    /* :695 */
        if (iter > nsteps) break;                          /* :696 */
        d1 = robust_weights_1b(res, rw, n);                /* :697-713 */
        if (cmad_out) cmad_out[iter - 1] = d1;
    }
    return 0;
}

/* ---------------------------------------------------------------------------
 * Window of one output point by the reference's slide predicate
 * (src/curvefit_regression.f90:663-669) evaluated as a binary search.  The
 * predicate d1 <= d2 is monotone in nleft for ascending x, so the first
 * nleft in [1, n-ns+1] that satisfies it (or n-ns+1) is what the sequential
 * walk reaches.  tests/test_oracle.py checks this against cf_oracle_lowess's
 * recorded windows.  i is 1-based; x is 1-based.
 * ------------------------------------------------------------------------- */
static int window_left_1b(const double *x, int n, int ns, int i)
{
    int lo = 1, hi = n - ns + 1, mid;
    double d1, d2;
    while (lo < hi) {
        mid = lo + (hi - lo) / 2;
        d1 = x[i] - x[mid];
        d2 = x[mid + ns] - x[i];          /* x(nright+1), nright = mid+ns-1 */
        if (d1 <= d2) hi = mid;
        else lo = mid + 1;
    }
    return lo;
}

/* ---------------------------------------------------------------------------
 * Teacher-forced evaluation: one lowest() call (src/curvefit_regression.f90:
 * 671-673) at each of m selected points idx[] (0-based), with the robustness
 * weights rw_in (NULL => userw = .false., i.e. the first pass).  Used to
 * check device passes at sizes where the full oracle cannot run.
 * ------------------------------------------------------------------------- */
ORACLE_API int cf_oracle_lowest_at(const double *x0, const double *y0, int n,
                                   double f, const double *rw_in,
                                   const int *idx, int m, double *ys_out,
                                   int *nleft_out, int *nright_out,
                                   int *nrt_out)
{
    const double *x = x0 - 1, *y = y0 - 1;
    const double *rw = rw_in ? rw_in - 1 : NULL;
    int ns = cf_oracle_span(n, f);
    int fail = 0;
    if (n < 2) return CF_INVALID_INPUT_ERROR;
#pragma omp parallel
    {
        double *w = (double *)malloc(sizeof(double) * (size_t)(n + 1));
        int k;
        if (!w) {
#pragma omp atomic write
            fail = 1;
        } else {
#pragma omp for schedule(dynamic, 1)
            for (k = 0; k < m; ++k) {
                int i = idx[k] + 1, nl, nr, nrt, ok;
                double v = 0.0;
                nl = window_left_1b(x, n, ns, i);
                nr = nl + ns - 1;
                ok = lowest_1b(x, y, n, x[i], &v, nl, nr, w, rw != NULL, rw,
                               &nrt);
                if (!ok) v = y[i];
                ys_out[k] = v;
                if (nleft_out) nleft_out[k] = nl;
                if (nright_out) nright_out[k] = nr;
                if (nrt_out) nrt_out[k] = nrt;
            }
            free(w);
        }
    }
    return fail ? CF_OUT_OF_MEMORY_ERROR : 0;
}

/* ---------------------------------------------------------------------------
 * Multi-threaded lowess with delta = 0 (what ls_smooth always passes,
 * src/curvefit_regression.f90:807,866).  Output points are split over
 * threads; each point's sums run in the reference's own order inside
 * lowest_1b, windows come from window_left_1b, tie groups are resolved as
 * :683-689 does (first member fitted, the others copy), so results are
 * bit-identical to cf_oracle_lowess(delta = 0) for ascending x.  Used as the all-cores CPU
 * baseline.  The scratch w is a per-thread window-sized buffer instead of the
 * reference's res alias; res is rebuilt at :695 either way.
 * ------------------------------------------------------------------------- */
ORACLE_API int cf_oracle_lowess_mt(const double *x0, const double *y0, int n,
                                   double f, int nsteps, double *ys0,
                                   double *rw0, double *res0, int nthreads,
                                   double *cmad_out)
{
    const double *x = x0 - 1, *y = y0 - 1;
    double *ys = ys0 - 1, *rw = rw0 - 1, *res = res0 - 1;
    int ns = cf_oracle_span(n, f), iter, fail = 0;
    if (n < 2) {
        int i;
        for (i = 1; i <= n; ++i) ys[i] = y[i];
        return 0;
    }
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
    for (iter = 1; iter <= nsteps + 1; ++iter) {
#pragma omp parallel
        {
            /* w is indexed with the global j; only [nleft, nrt] is touched */
            double *w = (double *)malloc(sizeof(double) * (size_t)(n + 1));
            int i;
            if (!w) {
#pragma omp atomic write
                fail = 1;
            } else {
#pragma omp for schedule(dynamic, 64)
                for (i = 1; i <= n; ++i) {
                    int nl, nr, nrt, ok;
                    if (i > 1 && x[i] == x[i - 1]) continue;  /* follower */
                    nl = window_left_1b(x, n, ns, i);
                    nr = nl + ns - 1;
                    ok = lowest_1b(x, y, n, x[i], &ys[i], nl, nr, w, iter > 1,
                                   rw, &nrt);
                    if (!ok) ys[i] = y[i];
                }
                free(w);
            }
        }
        if (fail) return CF_OUT_OF_MEMORY_ERROR;
        {
            /* :683-689 with delta = 0 and ascending x: cut = x(last), the
             * scan stops at the first x(i) > x(last), so the followers of a
             * fitted point are exactly the later points with an identical x
             * (their difference 0 < eps); each copies ys of its predecessor. */
            int i;
            for (i = 2; i <= n; ++i)
                if (x[i] == x[i - 1]) ys[i] = ys[i - 1];
        }
        {
            int i;
            for (i = 1; i <= n; ++i) res[i] = y[i] - ys[i];
        }
        if (iter > nsteps) break;
        {
            double c = robust_weights_1b(res, rw, n);
            if (cmad_out) cmad_out[iter - 1] = c;
        }
    }
    return 0;
}

/* ---------------------------------------------------------------------------
 * Index-tracking sort: LINALG/src/linalg_sorting.f90:310-354
 * (dbl_partition_ind, ascending branch) and :277-292 (qsort_dbl_ind).
 * x, ind 1-based slices of length n.
 * ------------------------------------------------------------------------- */
static int partition_ind_1b(double *x, int *ind, int n)
{
    int i, j, itemp;
    double temp, pivot;
    pivot = x[1];                                          /* :322 */
    i = 0;
    j = n + 1;
    for (;;) {
        j = j - 1;
        for (;;) { if (x[j] <= pivot) break; j = j - 1; }
        i = i + 1;
        for (;;) { if (x[i] >= pivot) break; i = i + 1; }
        if (i < j) {
            temp = x[i]; x[i] = x[j]; x[j] = temp;
            itemp = ind[i]; ind[i] = ind[j]; ind[j] = itemp;
        } else if (i == j) {
            return i + 1;
        } else {
            return i;
        }
    }
}

static void qsort_ind_1b(double *x, int *ind, int n)
{
    /* The reference recurses on both halves (:288-289); recursing on the
     * second half is a tail call, written here as a loop so that sorted
     * input (depth n in the reference) does not overflow the C stack.  The
     * sequence of partitions and swaps is unchanged. */
    while (n > 1) {
        int iq = partition_ind_1b(x, ind, n);
        qsort_ind_1b(x, ind, iq - 1);      /* x(:iq-1) */
        x += iq - 1;                       /* x(iq:)   */
        ind += iq - 1;
        n -= iq - 1;
    }
}

/* sort(x, ind, .true.): LINALG/src/linalg_sorting.f90:38-78.  ind holds
 * 1-based indices like the reference's tracking array. */
ORACLE_API void cf_oracle_sort_ind(double *x0, int *ind0, int n)
{
    if (n <= 1) return;                                    /* :74 */
    qsort_ind_1b(x0 - 1, ind0 - 1, n);
}

/* ---------------------------------------------------------------------------
 * ls_init data path: src/curvefit_regression.f90:791-804.
 * srt != 0: m_x = x; indices = 1..n; sort(m_x, indices); m_y = y(indices).
 * ------------------------------------------------------------------------- */
ORACLE_API int cf_oracle_ls_init(const double *x, const double *y, int n,
                                 int srt, double *m_x, double *m_y)
{
    int i;
    if (n < 0) return CF_INVALID_INPUT_ERROR;
    if (srt) {
        int *indices = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
        if (!indices) return CF_OUT_OF_MEMORY_ERROR;       /* :785-789 */
        for (i = 0; i < n; ++i) { m_x[i] = x[i]; indices[i] = i + 1; }
        cf_oracle_sort_ind(m_x, indices, n);               /* :797 */
        for (i = 0; i < n; ++i) m_y[i] = y[indices[i] - 1];/* :798 */
        free(indices);
    } else {
        for (i = 0; i < n; ++i) { m_x[i] = x[i]; m_y[i] = y[i]; }
    }
    return 0;
}

/* ls_smooth: src/curvefit_regression.f90:832-868 -- lowess with nsteps = 2,
 * delta = m_delta = 0 (:807, :866). */
ORACLE_API int cf_oracle_ls_smooth(const double *m_x, const double *m_y, int n,
                                   double f, double *ys, double *m_weights,
                                   double *m_residuals)
{
    return cf_oracle_lowess(m_x, m_y, n, f, 2, 0.0, ys, m_weights,
                            m_residuals, NULL, NULL, NULL, NULL);
}

/* The robustness step on its own (:697-713): |res| -> full ascending sort -> cmad ->
 * bisquare weights; single-threaded like the reference.  bench.py times it so that the
 * extrapolated CPU smooth() includes the sort. */
ORACLE_API double cf_oracle_robust_weights(const double *res0, double *rw0, int n)
{
    return robust_weights_1b(res0 - 1, rw0 - 1, n);
}

/* Number of OpenMP threads of the multi-threaded entry points from now on (they default
 * to OMP_NUM_THREADS, which launchers such as torchrun set to 1).  Returns the value in
 * effect. */
ORACLE_API int cf_oracle_set_threads(int nthreads)
{
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
    return omp_get_max_threads();
#else
    (void)nthreads;
    return 1;
#endif
}

ORACLE_API int cf_oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
