/*
 * lowess_c_abi.c -- the LOWESS path through the C ABI from plain C: the calls the
 * Fortran shim (fortran/curvefit_regression_cuda.f90) makes in place of ls_init's
 * sort and ls_smooth's lowess (src/curvefit_regression.f90:791-799, :866), on the data
 * of the reference's own example (examples/curvefit_lowess_example_2.f90: 100 points
 * of 0.5 sin(20 x) + cos(5 x) exp(-0.1 x) + noise on [0, 1]; the noise here is a
 * fixed LCG, the reference's is unseeded).
 *
 *   gcc -std=c99 -Iinclude examples/lowess_c_abi.c -Lcurvefit_b200 -lcurvefit_cuda \
 *       -Wl,-rpath,$PWD/curvefit_b200 -lm -o lowess_c_abi && ./lowess_c_abi
 *
 * Exit status: 0 on success, 3 when there is no CUDA device (the library has no CPU
 * fallback and says so), 1 on any other error.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "curvefit_cuda.h"

static int fail(const char* what, int32_t rc)
{
    char msg[256];
    cf_last_error(msg, (int32_t)sizeof msg);
    fprintf(stderr, "%s failed with code %d: %s\n", what, (int)rc, msg);
    return rc == CF_DEVICE_ERROR ? 3 : 1;
}

int main(void)
{
    enum { N = 100 };
    double x[N], y[N], xs[N], ys_in[N], ys[N], rw[N], res[N], ys2[N], rw2[N], res2[N];
    uint32_t lcg = 12345u;
    int32_t rc, i;

    /* shuffled abscissas, so that the sort has something to do */
    for (i = 0; i < N; ++i) {
        const int k = (int)((37u * (uint32_t)i + 11u) % (uint32_t)N);   /* a permutation: gcd(37, 100) = 1 */
        x[i] = (double)k / (double)(N - 1);
        lcg = 1664525u * lcg + 1013904223u;
        y[i] = 0.5 * sin(20.0 * x[i]) + cos(5.0 * x[i]) * exp(-0.1 * x[i])
             + ((double)(lcg >> 8) / 16777216.0 - 0.5);
    }

    printf("libcurvefit_cuda version %d, span for f = 0.2: %d points\n",
           (int)cf_version(), (int)cf_lowess_span(N, 0.2));

    rc = cf_sort_xy(x, y, N, xs, ys_in, NULL);            /* ls_init, srt = .true. */
    if (rc != CF_OK) return fail("cf_sort_xy", rc);
    rc = cf_lowess_smooth(xs, ys_in, N, 0.2, 2, ys, rw, res);   /* ls_smooth: nsteps = 2, delta = 0 */
    if (rc != CF_OK) return fail("cf_lowess_smooth", rc);

    for (i = 0; i < N; i += 11)
        printf("x = %.4f  y = % .4f  ys = % .4f  rw = %.4f  res = % .4f\n", xs[i], ys_in[i], ys[i], rw[i], res[i]);
    for (i = 0; i < N; ++i)
        if (!(res[i] == ys_in[i] - ys[i])) { fprintf(stderr, "res != y - ys at %d\n", (int)i); return 1; }

    /* what the shim's smooth() calls for long series: every visible GPU from this one process
     * (devices = NULL); a series this short runs on one device and must give the same answer */
    rc = cf_lowess_smooth_multi(xs, ys_in, N, 0.2, 2, NULL, 0, ys2, rw2, res2);
    if (rc != CF_OK) return fail("cf_lowess_smooth_multi", rc);
    for (i = 0; i < N; ++i)
        if (ys2[i] != ys[i] || rw2[i] != rw[i] || res2[i] != res[i]) {
            fprintf(stderr, "cf_lowess_smooth_multi differs at %d\n", (int)i);
            return 1;
        }
    cf_lowess_multi_release();
    return 0;
}
