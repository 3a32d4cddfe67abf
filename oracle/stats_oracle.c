/*
 * stats_oracle.c -- CPU restatement of the two routines next to the LOWESS path
 * (SURVEY 8f rows N3, N4).  TEST INFRASTRUCTURE ONLY, like lowess_oracle.c.
 *
 *   moving_average_1   src/curvefit_regression.f90:427-478
 *   mean_dbl           src/curvefit_statistics.f90:224-243   (running update)
 *   median_dbl         src/curvefit_statistics.f90:258-295   (sort = DLASRT 'I')
 *
 * Pinned by the reference's own known-answer test for median
 * (tests/curvefit_test_statistics.f90:286-309: {2,3,1,-4,-7,1,3} -> 1), see
 * tests/test_oracle.py.  moving_average has no test in the reference (only
 * examples/curvefit_moving_avg_example.f90:30 calls it): PARITY UNPINNED for it.
 */
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* mean_dbl, :224-243 */
static double mean_dbl(const double *x, int n)
{
    double z;
    int i;
    if (n == 0) return 0.0;                         /* :236-237 */
    z = x[0];                                       /* :239 */
    for (i = 2; i <= n; ++i)                        /* :240-242 */
        z = z + (x[i - 1] - z) / (double)i;
    return z;
}

/* moving_average_1, :427-478.  Returns 0, or 201 (CF_INVALID_INPUT_ERROR, :452-457). */
ORACLE_API int cf_oracle_moving_average(double *x, int n, int npts)
{
    double *buffer;
    int i;
    if (npts < 2 || npts > n) return 201;
    buffer = (double *)calloc((size_t)npts, sizeof(double));      /* :460, :468 */
    if (!buffer) return 105;
    for (i = 0; i < n; ++i) {                                      /* :469 */
        memmove(buffer + 1, buffer, sizeof(double) * (size_t)(npts - 1));   /* :471 */
        buffer[0] = x[i];                                          /* :474 */
        x[i] = mean_dbl(buffer, npts);                             /* :477 */
    }
    free(buffer);
    return 0;
}

static int cmp_dbl(const void *a, const void *b)
{
    const double u = *(const double *)a, v = *(const double *)b;
    return (u > v) - (u < v);
}

/* median_dbl, :258-295; srt != 0 sorts x in place (ascending values are algorithm
 * independent, so qsort stands in for DLASRT). */
ORACLE_API double cf_oracle_median(double *x, int n, int srt)
{
    const int nmid = n / 2, nmidp1 = nmid + 1, iflag = n - 2 * nmid;   /* :273-275 */
    if (n == 0) return 0.0;                                            /* :280-283 */
    if (srt) qsort(x, (size_t)n, sizeof(double), cmp_dbl);             /* :286 */
    if (iflag == 0) return 0.5 * (x[nmid - 1] + x[nmidp1 - 1]);        /* :289-290 */
    return x[nmidp1 - 1];                                              /* :292 */
}
