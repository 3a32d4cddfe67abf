// windows.cu -- neighbourhood search for every output point.
//
// Replaces the sequential slide loop of lowess (src/curvefit_regression.f90:
// 658-669) and the r <= h9 / r <= h1 / exit tests of lowest's first sweep
// (:542-556) with independent binary searches per point.  The slide predicate
// d1 <= d2 is monotone in nleft and in i for ascending x (fl(a-b) is monotone
// in both arguments), so the first nleft that satisfies it is what the
// sequential walk reaches; every comparison below uses the reference's own
// floating-point expressions (no algebraic rewriting, no FMA), which makes
// all outputs integer-exact.
#include "common.cuh"

namespace cf {

__device__ __forceinline__ void windows_point(const double* __restrict__ x, int32_t n, int32_t ns,
                                              Windows w, int32_t i0, int32_t& out_l9, int32_t& out_r9);

__global__ void __launch_bounds__(256)
windows_kernel(const double* __restrict__ x, int32_t n, int32_t ns, Windows w, int32_t begin,
               int32_t end, int32_t* __restrict__ span, const int32_t* __restrict__ fidx)
{
    // only the handle's shard [begin, end) of output points is searched; span[0..1]
    // receives min lo9 / max hi9 over it (the elements its fits can touch).  delta > 0
    // (delta.cu): slots [begin, end) of the fitted-point list fidx instead.
    const int32_t s0 = begin + blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = s0 < end;
    int32_t l9 = INT32_MAX, r9 = -1;
    if (live) windows_point(x, n, ns, w, fidx ? fidx[s0] : s0, l9, r9);
    l9 = __reduce_min_sync(0xffffffffu, l9);
    r9 = __reduce_max_sync(0xffffffffu, r9);
    if ((threadIdx.x & 31) == 0 && r9 >= 0) {
        int32_t* sp = span + 2 * blockIdx.y;
        atomicMin(&sp[0], l9);
        atomicMax(&sp[1], r9);
    }
}


__device__ __forceinline__ void windows_point(const double* __restrict__ x, int32_t n, int32_t ns,
                                              Windows w, int32_t i0, int32_t& out_l9, int32_t& out_r9)
{
    // batched mode: series blockIdx.y lives at offset blockIdx.y * n in every array;
    // all indices stay relative to the series
    const size_t off = (size_t)blockIdx.y * (size_t)n;
    x += off;
    w.nleft += off; w.lo9 += off; w.hi9 += off; w.lo1 += off; w.hi1 += off; w.leader += off;

    // tie group leader: first index whose x equals x[i0] (:683-689 -- with
    // delta = 0 the followers of a fitted point are the later identical x).
    const double xi = x[i0];
    int32_t i = i0;
    if (i0 > 0 && x[i0 - 1] == xi) {
        int32_t lo = 0, hi = i0 - 1;  // x[hi] == xi
        while (lo < hi) {
            const int32_t mid = lo + ((hi - lo) >> 1);
            if (x[mid] < xi) lo = mid + 1; else hi = mid;
        }
        i = lo;
    }
    const double xs = x[i];

    // nleft: first L in [0, n-ns] with x(i)-x(L) <= x(L+ns)-x(i)   (:664-666)
    int32_t lo = 0, hi = n - ns;
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        const double d1 = __dsub_rn(xs, x[mid]);
        const double d2 = __dsub_rn(x[mid + ns], xs);
        if (d1 <= d2) hi = mid; else lo = mid + 1;
    }
    const int32_t nl = lo, nr = lo + ns - 1;

    // bandwidth (:536-538)
    const double hl = __dsub_rn(xs, x[nl]);
    const double hr = __dsub_rn(x[nr], xs);
    const double h = hl > hr ? hl : hr;
    const double h9 = __dmul_rn(0.999, h);
    const double h1 = __dmul_rn(0.001, h);

    // support [lo9, hi9]: |x_j - xs| <= h9, j >= nleft (:542-556)
    int32_t a = nl, b = i;  // predicate true at j = i (r = 0)
    while (a < b) {
        const int32_t mid = a + ((b - a) >> 1);
        if (fabs(__dsub_rn(x[mid], xs)) <= h9) b = mid; else a = mid + 1;
    }
    const int32_t l9 = a;
    a = i; b = n - 1;
    while (a < b) {
        const int32_t mid = a + ((b - a + 1) >> 1);
        if (fabs(__dsub_rn(x[mid], xs)) <= h9) a = mid; else b = mid - 1;
    }
    const int32_t r9 = a;

    // core [lo1, hi1]: |x_j - xs| <= h1 inside the support (:547-551, w = 1)
    a = l9; b = i;
    while (a < b) {
        const int32_t mid = a + ((b - a) >> 1);
        if (fabs(__dsub_rn(x[mid], xs)) <= h1) b = mid; else a = mid + 1;
    }
    const int32_t l1 = a;
    a = i; b = r9;
    while (a < b) {
        const int32_t mid = a + ((b - a + 1) >> 1);
        if (fabs(__dsub_rn(x[mid], xs)) <= h1) a = mid; else b = mid - 1;
    }
    const int32_t r1 = a;

    w.nleft[i0] = nl;
    w.lo9[i0] = l9;
    w.hi9[i0] = r9;
    w.lo1[i0] = l1;
    w.hi1[i0] = r1;
    w.leader[i0] = i;
    out_l9 = l9;
    out_r9 = r9;
}

__global__ void span_init_kernel(int32_t* span, int32_t nseries)
{
    const int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nseries) { span[2 * s] = INT32_MAX; span[2 * s + 1] = -1; }
}

int launch_windows(cf_lowess* h, int32_t ns)
{
    if (h->n < 2) return CF_OK;
    const int threads = 256;
    const bool listed = h->delta > 0.0;
    const int32_t begin = listed ? 0 : h->begin, end = listed ? h->nfit : h->end;
    const int cnt = end - begin;
    span_init_kernel<<<(h->nseries + 255) / 256, 256, 0, h->stream>>>(h->span, h->nseries);
    if (cnt <= 0) return CF_OK;
    const int blocks = (cnt + threads - 1) / threads;
    windows_kernel<<<dim3(blocks, h->nseries), threads, 0, h->stream>>>(h->x, h->n, ns, h->win, begin, end,
                                                                        h->span, listed ? h->fidx : nullptr);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

}  // namespace cf
