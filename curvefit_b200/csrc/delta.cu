// delta.cu -- the delta > 0 mode of lowess (SURVEY 8f row N1;
// src/curvefit_regression.f90:674-690): lowest() is evaluated only at points
// spaced more than delta apart, points within delta of the last fitted point are
// linearly interpolated, points closer than eps to it copy its value.  The
// reference's own smooth() never reaches this code (m_delta = 0, :807); parity is
// defined by the restated oracle, which takes delta like the reference's lowess.
//
// The reference selects the fitted points by a sequential greedy walk.  Here:
//   next(p)  = the point fitted after p                  (one thread per p)
//            = max(c(p) + 1, i(p) - 1),  c(p) = end of the chain of points closer
//              than eps to their predecessor starting at p (`last` after :683-689),
//              i(p) = first index with x > x(p) + delta (:682-684); n if c(p) is
//              the last point (:692).
//   next is strictly increasing past p and monotone in p, hence the fitted set
//   F = {0, next(0), next(next(0)), ...} satisfies, with D(q) = |F below q|,
//       D(q) = D(par(q)) + 1,   par(q) = the smallest p with next(p) >= q,
//   i.e. D is the depth of q in a tree rooted at 0 -- computed for all q at once by
//   pointer doubling (log2 |F| rounds), and p is fitted iff D(p + 1) > D(p).  Its
//   position in the ascending list of fitted points is D(p) itself.
// Everything is integer / comparison work on x with the reference's own
// floating-point expressions, so F is exactly the reference's.
#include <float.h>

#include "common.cuh"

namespace cf {

__global__ void __launch_bounds__(256)
delta_next_kernel(const double* __restrict__ x, int32_t n, double delta,
                  int32_t* __restrict__ nxt, int32_t* __restrict__ cend)
{
    const int32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > n) return;
    if (p == n) { nxt[n] = n; return; }
    const double xp = x[p];
    const double cut = __dadd_rn(xp, delta);                      // :682
    // i(p): first index in (p, n] with x > cut (:683-684)
    int32_t lo = p + 1, hi = n;
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (x[mid] > cut) hi = mid; else lo = mid + 1;
    }
    const int32_t ip = lo;
    // c(p): the run of values equal to x(p) ...
    int32_t a = p, b = ip - 1;
    while (a < b) {
        const int32_t mid = a + ((b - a + 1) >> 1);
        if (x[mid] == xp) a = mid; else b = mid - 1;
    }
    int32_t c = a;
    // ... continued while consecutive points differ by less than eps (:685-688)
    while (c + 1 < ip && fabs(__dsub_rn(x[c + 1], x[c])) < DBL_EPSILON) ++c;
    cend[p] = c;
    nxt[p] = c >= n - 1 ? n : max(c + 1, ip - 1);                 // :690, :692
}

// par(q) = smallest p with next(p) >= q, q in [1, n]; depth 1.  Root: q = 0.
__global__ void __launch_bounds__(256)
delta_parent_kernel(const int32_t* __restrict__ nxt, int32_t n, int32_t* __restrict__ par,
                    int32_t* __restrict__ dep)
{
    const int32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q > n) return;
    if (q == 0) { par[0] = 0; dep[0] = 0; return; }
    int32_t lo = 0, hi = q - 1;              // next(q - 1) >= q always
    while (lo < hi) {
        const int32_t mid = lo + ((hi - lo) >> 1);
        if (nxt[mid] >= q) hi = mid; else lo = mid + 1;
    }
    par[q] = lo;
    dep[q] = 1;
}

__global__ void __launch_bounds__(256)
delta_jump_kernel(const int32_t* __restrict__ par, const int32_t* __restrict__ dep, int32_t n,
                  int32_t* __restrict__ par2, int32_t* __restrict__ dep2, int32_t* __restrict__ busy)
{
    const int32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q > n) return;
    const int32_t p = par[q];
    const int32_t pp = par[p];
    par2[q] = pp;
    dep2[q] = dep[q] + dep[p];
    if (pp != 0) *busy = 1;
}

// fitted points in ascending order with the end of their tie chain
__global__ void __launch_bounds__(256)
delta_list_kernel(const int32_t* __restrict__ D, const int32_t* __restrict__ cend, int32_t n,
                  int32_t* __restrict__ fidx, int32_t* __restrict__ fcend)
{
    const int32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const int32_t d = D[p];
    if (D[p + 1] > d) {
        fidx[d] = p;
        fcend[d] = cend[p];
    }
}

// Points that were not fitted (:674-680, :685-688), then res = y - ys (:695).
__global__ void __launch_bounds__(256)
delta_fill_kernel(const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ ys,
                  double* __restrict__ res, int32_t begin, int32_t end, const int32_t* __restrict__ D,
                  const int32_t* __restrict__ fidx, const int32_t* __restrict__ fcend)
{
    const int32_t j = begin + blockIdx.x * blockDim.x + threadIdx.x;     // the shard's points
    if (j >= end) return;
    const int32_t k = D[j + 1] - 1;          // the last fitted point at or before j
    const int32_t a = fidx[k];
    double v = ys[a];
    if (j != a) {
        const int32_t c = fcend[k];          // `last` when the next point is fitted
        if (j > c) {
            const int32_t b = fidx[k + 1];
            const double xl = x[c];
            const double denom = __dsub_rn(x[b], xl);                          // :675
            const double alpha = __ddiv_rn(__dsub_rn(x[j], xl), denom);        // :677
            v = __dadd_rn(__dmul_rn(alpha, ys[b]), __dmul_rn(__dsub_rn(1.0, alpha), v));   // :678
        }
        ys[j] = v;
    }
    res[j] = __dsub_rn(y[j], v);
}

void delta_free(cf_lowess* h)
{
    cudaFree(h->fidx); cudaFree(h->fcend); cudaFree(h->frank);
    h->fidx = h->fcend = h->frank = nullptr;
    h->cap_delta = 0;
    h->delta_built = -1.0;
    h->nfit = 0;
}

int delta_prepare(cf_lowess* h)
{
    if (!(h->delta > 0.0) || h->delta_built == h->delta) return CF_OK;
    if (h->nseries != 1) {
        set_error("delta > 0 is supported for a single series (not for batched handles)");
        return CF_INVALID_INPUT_ERROR;
    }
    if (h->x_partial) {
        // after a span upload x is -inf / +inf outside the span; the list built from that (:682-690)
        // would differ from the reference's and from rank to rank
        set_error("delta > 0 needs the whole of x on the handle (cf_lowess_init), not a span upload "
                  "(cf_lowess_init_span)");
        return CF_INVALID_INPUT_ERROR;
    }
    const int32_t n = h->n;
    const size_t len = (size_t)n + 1;
    if ((int64_t)len > h->cap_delta) {
        delta_free(h);
        CF_CUDA(cudaMalloc(&h->fidx, sizeof(int32_t) * len));
        CF_CUDA(cudaMalloc(&h->fcend, sizeof(int32_t) * len));
        CF_CUDA(cudaMalloc(&h->frank, sizeof(int32_t) * len));
        h->cap_delta = (int64_t)len;
    }
    // scratch: next, chain ends, two (parent, depth) pairs, busy flag
    int32_t* tmp = nullptr;
    CF_CUDA(cudaMalloc(&tmp, sizeof(int32_t) * (6 * len + 1)));
    int32_t *nxt = tmp, *cend = tmp + len, *par[2] = {tmp + 2 * len, tmp + 3 * len},
            *dep[2] = {tmp + 4 * len, tmp + 5 * len}, *busy = tmp + 6 * len;
    const int blocks = (int)((len + 255) / 256);
    cudaStream_t s = h->stream;
    int rc = CF_OK;
    auto fail = [&](cudaError_t e, const char* what) { rc = cuda_fail(e, what, __FILE__, __LINE__); };
    delta_next_kernel<<<blocks, 256, 0, s>>>(h->x, n, h->delta, nxt, cend);
    delta_parent_kernel<<<blocks, 256, 0, s>>>(nxt, n, par[0], dep[0]);
    h->launches += 2;
    int cur = 0;
    for (int round = 0; round < 32 && rc == CF_OK; ++round) {
        int flag = 0;
        cudaError_t e = cudaMemsetAsync(busy, 0, sizeof(int32_t), s);
        if (e != cudaSuccess) { fail(e, "delta memset"); break; }
        delta_jump_kernel<<<blocks, 256, 0, s>>>(par[cur], dep[cur], n, par[cur ^ 1], dep[cur ^ 1], busy);
        h->launches++;
        cur ^= 1;
        e = cudaMemcpyAsync(&flag, busy, sizeof(int32_t), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { fail(e, "delta doubling"); break; }
        if (!flag) break;
    }
    if (rc == CF_OK) {
        // every parent is the root now: dep = D
        cudaError_t e = cudaMemcpyAsync(h->frank, dep[cur], sizeof(int32_t) * len, cudaMemcpyDeviceToDevice, s);
        if (e != cudaSuccess) fail(e, "delta rank copy");
    }
    if (rc == CF_OK) {
        delta_list_kernel<<<blocks, 256, 0, s>>>(h->frank, cend, n, h->fidx, h->fcend);
        h->launches++;
        int32_t m = 0;
        cudaError_t e = cudaMemcpyAsync(&m, h->frank + n, sizeof(int32_t), cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) fail(e, "delta list");
        else { h->nfit = m; h->delta_built = h->delta; }
    }
    if (rc == CF_OK) {
        // Range-sharded series: every rank derives the whole list (it holds all of x) and fits the
        // slots of its own block [begin, end) plus the last fitted point at or before `begin` and
        // the first one past the block -- delta_fill_kernel interpolates between them (:675-679) --
        // so no fitted value has to cross ranks.  D(p) = number of fitted points below p.
        h->slot_begin = 0;
        h->slot_end = h->nfit;
        if (h->nranks > 1 && h->end > h->begin) {
            int32_t d[2] = {0, 0};
            cudaError_t e = cudaMemcpyAsync(&d[0], h->frank + h->begin + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaMemcpyAsync(&d[1], h->frank + h->end, sizeof(int32_t), cudaMemcpyDeviceToHost, s);
            if (e == cudaSuccess) e = cudaStreamSynchronize(s);
            if (e != cudaSuccess) fail(e, "delta slots");
            else {
                h->slot_begin = d[0] > 0 ? d[0] - 1 : 0;
                h->slot_end = d[1] + 1 < h->nfit ? d[1] + 1 : h->nfit;
            }
        } else if (h->nranks > 1) {
            h->slot_end = 0;                               // empty shard
        }
    }
    cudaFree(tmp);
    return rc;
}

int launch_delta_fill(cf_lowess* h)
{
    const int32_t cnt = h->end - h->begin;
    if (cnt <= 0) return CF_OK;
    delta_fill_kernel<<<(cnt + 255) / 256, 256, 0, h->stream>>>(h->x, h->y, h->ys, h->res, h->begin, h->end,
                                                                h->frank, h->fidx, h->fcend);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

}  // namespace cf
