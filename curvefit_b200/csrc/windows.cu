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

// Guided searches.  Every search below is "first index of [lo, hi] where a monotone
// predicate turns true", so WHICH probes find it does not change the answer; a plain
// bisection is 22-27 DEPENDENT loads at L2 latency per search and the kernel was bound by
// exactly that (n = 1e8: 7.9 ms, 360 cycles per step).  first_true() starts from a guess g,
// gallops away from it in doubling steps until the answer is bracketed and bisects the
// bracket: 2 probes when the guess is right, 2 log2(error) otherwise, never more than
// twice a bisection.  Guesses: on the first point of a thread the uniform-grid positions
// (nleft = i - ns/2, band edges by linear interpolation); on its later points the previous
// point's answers shifted by the distance between the two points.
template <class P>
__device__ __forceinline__ int32_t first_true(int32_t lo, int32_t hi, int32_t g, P pred)
{
    // contract: pred(hi) holds by construction and is never evaluated
    g = min(max(g, lo), hi);
    int32_t a, b;
    if (g == hi || pred(g)) {                      // answer <= g: gallop down
        a = lo; b = g;
        int32_t step = 1;
        while (b > lo) {
            const int32_t t = (b - lo < step) ? lo : b - step;
            if (pred(t)) { b = t; step <<= 1; } else { a = t + 1; break; }
        }
        if (b == lo) a = lo;
    } else {                                       // answer > g: gallop up
        a = g + 1; b = hi;
        int32_t step = 1;
        while (a < hi) {
            if (hi - a < step) break;              // the bracket is [a, hi]
            const int32_t t = a + step - 1;
            if (pred(t)) { b = t; break; }
            a = t + 1; step <<= 1;
        }
    }
    while (a < b) {
        const int32_t mid = a + ((b - a) >> 1);
        if (pred(mid)) b = mid; else a = mid + 1;
    }
    return a;
}

// last index of [lo, hi] where a monotone (true .. false) predicate holds; pred(lo) holds by
// construction and is never evaluated
template <class P>
__device__ __forceinline__ int32_t last_true(int32_t lo, int32_t hi, int32_t g, P pred)
{
    return first_true(lo, hi, g, [&](int32_t j) { return !pred(j + 1); });
}

// hints carried from one point of a thread to its next
struct WinHint {
    int32_t i, nl, l9, r9, l1, r1;     // i < 0: none yet
};

__device__ __forceinline__ void windows_point(const double* __restrict__ x, int32_t n, int32_t ns,
                                              Windows w, int32_t i0, WinHint& hint);


__global__ void __launch_bounds__(256)
windows_kernel(const double* __restrict__ x, int32_t n, int32_t ns, Windows w, int32_t begin,
               int32_t end, int32_t* __restrict__ span, const int32_t* __restrict__ fidx, int per_thread)
{
    // only the handle's shard [begin, end) of output points is searched; span[0..1]
    // receives min lo9 / max hi9 over it (the elements its fits can touch).  delta > 0
    // (delta.cu): slots [begin, end) of the fitted-point list fidx instead.
    // A warp takes 32 * per_thread consecutive slots, lane l the slots l, l + 32, ...
    // (coalesced stores; each search starts from the lane's previous answers).
    const int32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int32_t lane = threadIdx.x & 31;
    const int64_t first = (int64_t)begin + (int64_t)warp * 32 * per_thread + lane;
    WinHint hint = {-1, 0, 0, 0, 0, 0};
    int32_t l9 = INT32_MAX, r9 = -1;
    for (int k = 0; k < per_thread; ++k) {
        const int64_t s0 = first + 32 * k;
        if (s0 >= end) break;
        windows_point(x, n, ns, w, fidx ? fidx[s0] : (int32_t)s0, hint);
        l9 = min(l9, hint.l9);
        r9 = max(r9, hint.r9);
    }
    l9 = __reduce_min_sync(0xffffffffu, l9);
    r9 = __reduce_max_sync(0xffffffffu, r9);
    if (lane == 0 && r9 >= 0) {
        int32_t* sp = span + 2 * blockIdx.y;
        atomicMin(&sp[0], l9);
        atomicMax(&sp[1], r9);
    }
}

// position of value v between (ja, xa) and (jb, xb) by linear interpolation, as a guess
__device__ __forceinline__ int32_t guess_between(int32_t ja, double xa, int32_t jb, double xb, double v)
{
    const double den = xb - xa;
    if (!(den > 0.0)) return ja;
    double t = (v - xa) / den;
    t = fmin(fmax(t, 0.0), 1.0);
    return ja + (int32_t)(t * (double)(jb - ja));
}

__device__ __forceinline__ void windows_point(const double* __restrict__ x, int32_t n, int32_t ns,
                                              Windows w, int32_t i0, WinHint& hint)
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
    const bool hinted = hint.i >= 0;
    const int32_t shift = i - hint.i;

    // nleft: first L in [0, n-ns] with x(i)-x(L) <= x(L+ns)-x(i)   (:664-666)
    const int32_t nl = first_true(0, n - ns, hinted ? hint.nl + shift : i - ns / 2, [&](int32_t L) {
        const double d1 = __dsub_rn(xs, x[L]);
        const double d2 = __dsub_rn(x[L + ns], xs);
        return d1 <= d2;
    });
    const int32_t nr = nl + ns - 1;

    // bandwidth (:536-538)
    const double xl = x[nl], xr = x[nr];
    const double hl = __dsub_rn(xs, xl);
    const double hr = __dsub_rn(xr, xs);
    const double h = hl > hr ? hl : hr;
    const double h9 = __dmul_rn(0.999, h);
    const double h1 = __dmul_rn(0.001, h);
    auto in9 = [&](int32_t j) { return fabs(__dsub_rn(x[j], xs)) <= h9; };
    auto in1 = [&](int32_t j) { return fabs(__dsub_rn(x[j], xs)) <= h1; };

    // support [lo9, hi9]: |x_j - xs| <= h9, j >= nleft (:542-556); true at j = i (r = 0)
    const int32_t l9 = first_true(nl, i, hinted ? hint.l9 + shift : guess_between(nl, xl, i, xs, xs - h9), in9);
    const int32_t r9 = last_true(i, n - 1, hinted ? hint.r9 + shift : guess_between(i, xs, nr, xr, xs + h9), in9);
    // core [lo1, hi1]: |x_j - xs| <= h1 inside the support (:547-551, w = 1)
    const int32_t l1 = first_true(l9, i, hinted ? hint.l1 + shift : guess_between(nl, xl, i, xs, xs - h1), in1);
    const int32_t r1 = last_true(i, r9, hinted ? hint.r1 + shift : guess_between(i, xs, nr, xr, xs + h1), in1);

    w.nleft[i0] = nl;
    w.lo9[i0] = l9;
    w.hi9[i0] = r9;
    w.lo1[i0] = l1;
    w.hi1[i0] = r1;
    w.leader[i0] = i;
    hint.i = i; hint.nl = nl; hint.l9 = l9; hint.r9 = r9; hint.l1 = l1; hint.r1 = r1;
}

__global__ void span_init_kernel(int32_t* span, int32_t nseries)
{
    const int32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nseries) { span[2 * s] = INT32_MAX; span[2 * s + 1] = -1; }
}

int launch_windows_range(cf_lowess* h, int32_t ns, int32_t begin, int32_t end, bool init_span)
{
    if (h->n < 2) return CF_OK;
    const int threads = 256;
    const bool listed = h->delta > 0.0;
    const int cnt = end - begin;
    if (init_span) span_init_kernel<<<(h->nseries + 255) / 256, 256, 0, h->stream>>>(h->span, h->nseries);
    if (cnt <= 0) return CF_OK;
    // points per thread: as many as leave about two waves of threads on the device
    const int64_t total = (int64_t)cnt * h->nseries;
    int per_thread = (int)(total / (2 * 148 * 2048));
    per_thread = per_thread < 1 ? 1 : per_thread > 8 ? 8 : per_thread;
    const int64_t nthreads = ((int64_t)cnt + 32 * per_thread - 1) / (32 * per_thread) * 32;
    const int blocks = (int)((nthreads + threads - 1) / threads);
    windows_kernel<<<dim3(blocks, h->nseries), threads, 0, h->stream>>>(h->x, h->n, ns, h->win, begin, end,
                                                                        h->span, listed ? h->fidx : nullptr, per_thread);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_windows(cf_lowess* h, int32_t ns)
{
    const bool listed = h->delta > 0.0;
    return launch_windows_range(h, ns, listed ? h->slot_begin : h->begin, listed ? h->slot_end : h->end, true);
}

}  // namespace cf
