// robust.cu -- the robustness step of lowess:
//   residuals + tie copy      src/curvefit_regression.f90:683-689, :695
//   median of |res|           :697-703  (full DLASRT sort in the reference;
//                             here an exact MSD radix SELECT of the two order
//                             statistics rw(m1), rw(m2) -- sorted values are
//                             algorithm independent, so the result is bit-equal)
//   bisquare weights          :704-713
// All kernels work on the handle's shard [begin, end) of the sorted index range;
// the digit histograms are the only data that must be summed across GPUs.
#include "common.cuh"
#include "peer.cuh"

namespace cf {

// ---- residuals ------------------------------------------------------------------
__global__ void __launch_bounds__(256)
residual_kernel(const double* __restrict__ y, double* __restrict__ ys,
                double* __restrict__ res, const int32_t* __restrict__ leader,
                int32_t begin, int32_t end, int32_t n)
{
    const int32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const size_t off = (size_t)blockIdx.y * (size_t)n;     // batched: series offset
    y += off; ys += off; res += off; leader += off;
    // a follower of a tie group takes the fitted value of the group's first
    // member (:685-688).  If that member belongs to the previous shard, the
    // shard's first point was fitted in its place (fit.cu, `valid`).
    const int32_t l = max(leader[i], begin);
    const double v = ys[l];
    if (l != i) ys[i] = v;
    res[i] = __dsub_rn(y[i], v);                                            // :695
}

int launch_residuals(cf_lowess* h)
{
    if (h->delta > 0.0) return launch_delta_fill(h);   // interpolation + tie copy + residuals
    const int cnt = h->end - h->begin;
    if (cnt <= 0) return CF_OK;
    residual_kernel<<<dim3((cnt + 255) / 256, h->nseries), 256, 0, h->stream>>>(
        h->y, h->ys, h->res, h->win.leader, h->begin, h->end, h->n);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

// ---- exact select of the two middle order statistics of |res| ------------------------
__device__ __forceinline__ int sel_shift(int round)
{
    const int s = 64 - kSelBits * (round + 1);
    return s > 0 ? s : 0;
}
__device__ __forceinline__ int sel_width(int round)
{
    const int s = 64 - kSelBits * (round + 1);
    return s >= 0 ? kSelBits : kSelBits + s;
}

__global__ void select_reset_kernel(SelectState* st, unsigned long long* hist, int32_t n)
{
    st += blockIdx.x;
    hist += (size_t)blockIdx.x * kHistLen;
    for (int k = threadIdx.x; k < kHistLen; k += blockDim.x) hist[k] = 0ull;
    if (threadIdx.x == 0) {
        st->prefix[0] = st->prefix[1] = 0ull;
        // 1-based m2 = n - m1 + 1 <= m1 = 1 + n/2  ->  0-based (n-1)/2 and n/2   (:699-700)
        st->rank[0] = (unsigned long long)((n - 1) / 2);
        st->rank[1] = (unsigned long long)(n / 2);
        // cmad, c1, c9 keep the previous iteration's values until pick_kernel's last round
    }
}

int launch_select_reset(cf_lowess* h)
{
    select_reset_kernel<<<h->nseries, 256, 0, h->stream>>>(h->sel, h->hist, h->n);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

__global__ void __launch_bounds__(512)
hist_kernel(const double* __restrict__ res, int32_t begin, int32_t end, int32_t n, int round,
            const SelectState* __restrict__ st, unsigned long long* __restrict__ hist, int ordered)
{
    __shared__ unsigned int sh[kHistLen];
    res += (size_t)blockIdx.y * (size_t)n;
    st += blockIdx.y;
    hist += (size_t)blockIdx.y * kHistLen;
    for (int k = threadIdx.x; k < kHistLen; k += blockDim.x) sh[k] = 0u;
    __syncthreads();
    const int shift = sel_shift(round), width = sel_width(round);
    const unsigned mask = (1u << width) - 1u;
    const unsigned long long p0 = st->prefix[0], p1 = st->prefix[1];
    const int hs = shift + width;  // bits below the resolved prefix
    const int32_t stride = gridDim.x * blockDim.x;
    for (int32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
        // ordered = 0: |res| (:697); 1: order-preserving image of the signed value (stats.cu)
        const unsigned long long key = ordered ? ordered_key(res[i])
            : (unsigned long long)__double_as_longlong(fabs(res[i]));
        const unsigned long long hi = hs >= 64 ? 0ull : (key >> hs);
        const unsigned d = (unsigned)(key >> shift) & mask;
        if (hi == p0) atomicAdd(&sh[d], 1u);
        if (hi == p1) atomicAdd(&sh[kSelBins + d], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < kHistLen; k += blockDim.x)
        if (sh[k]) atomicAdd(&hist[k], (unsigned long long)sh[k]);
}

int launch_hist(cf_lowess* h, int round)
{
    const int cnt = h->end - h->begin;
    if (cnt <= 0) return CF_OK;
    int blocks = (cnt + 512 * 8 - 1) / (512 * 8);
    if (blocks > 148 * 4) blocks = 148 * 4;
    hist_kernel<<<dim3(blocks, h->nseries), 512, 0, h->stream>>>(h->res, h->begin, h->end, h->n,
                                                                 round, h->sel, h->hist, 0);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}


// One warp per tracked rank: locate the bin holding the rank, extend the prefix,
// rebase the rank, clear the histogram for the next round; after the last round
// form cmad, c1, c9 (:701-703).
// PEER: the histogram is first summed over the ranks through peer memory (peer.cuh).
template <bool PEER>
__global__ void __launch_bounds__(PEER ? 256 : 64)
pick_kernel(SelectState* st, unsigned long long* hist, int round, const PeerPtrs pp,
            unsigned long long epoch)
{
    st += blockIdx.x;
    hist += (size_t)blockIdx.x * kHistLen;
    if (PEER) peer_hist_allreduce(pp, hist, epoch);
    const int t = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (t < 2) {
    const int width = sel_width(round);
    const int bins = 1 << width;
    const int per = kSelBins / 32;
    unsigned long long* hh = hist + t * kSelBins;
    unsigned long long sum = 0;
    for (int k = 0; k < per; ++k) {
        const int b = lane * per + k;
        sum += b < bins ? hh[b] : 0ull;
    }
    unsigned long long inc = sum;   // inclusive scan over lanes
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    const unsigned long long exc = inc - sum;
    const unsigned long long r = st->rank[t];
    const bool mine = r >= exc && r < inc;
    __syncwarp();
    if (mine) {
        unsigned long long cum = exc;
        for (int k = 0; k < per; ++k) {
            const int b = lane * per + k;
            const unsigned long long cnt = b < bins ? hh[b] : 0ull;
            if (r < cum + cnt) {
                st->prefix[t] = (st->prefix[t] << width) | (unsigned long long)b;
                st->rank[t] = r - cum;
                break;
            }
            cum += cnt;
        }
    }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < kHistLen; k += blockDim.x) hist[k] = 0ull;
    if (round == kSelRounds - 1 && threadIdx.x == 0) {
        const double lo = __longlong_as_double((long long)st->prefix[0]);   // rw(m2)
        const double hi = __longlong_as_double((long long)st->prefix[1]);   // rw(m1)
        const double cmad = __dmul_rn(3.0, __dadd_rn(hi, lo));              // :701
        st->cmad = cmad;
        st->c9 = __dmul_rn(0.999, cmad);                                    // :702
        st->c1 = __dmul_rn(0.001, cmad);                                    // :703
    }
}

int launch_pick(cf_lowess* h, int round)
{
    if (h->peer_attached)
        pick_kernel<true><<<1, 256, 0, h->stream>>>(h->sel, h->hist, round, h->peer, ++h->epoch_hist);
    else
        pick_kernel<false><<<h->nseries, 64, 0, h->stream>>>(h->sel, h->hist, round, PeerPtrs{}, 0ull);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

// ---- bisquare robustness weights (:704-713) ---------------------------------------------
// PEER: the weight is also stored into the rw vector of every rank that needs it.
template <bool PEER>
__global__ void __launch_bounds__(256)
bisquare_kernel(const double* __restrict__ res, double* __restrict__ rw,
                const SelectState* __restrict__ st, int32_t begin, int32_t end, int32_t n,
                const PeerPtrs pp)
{
    const int32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const size_t off = (size_t)blockIdx.y * (size_t)n;
    res += off; rw += off; st += blockIdx.y;
    const double cmad = st->cmad, c1 = st->c1, c9 = st->c9;
    const double w = bisquare_weight(res[i], cmad, c1, c9);
    rw[i] = w;
    if (PEER) {
        // halo exchange: only ranks whose windows reach element i get it
        const PeerComm* own = pp.comm[pp.rank];
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
            if (p < pp.nranks && p != pp.rank && i >= own->peer_span[p][0] && i <= own->peer_span[p][1])
                pp.rw[p][i] = w;
    }
}

int launch_bisquare(cf_lowess* h)
{
    const int cnt = h->end - h->begin;
    if (cnt <= 0) return launch_peer_barrier(h);   // an empty shard still meets the others
    if (h->peer_attached) {
        bisquare_kernel<true><<<(cnt + 255) / 256, 256, 0, h->stream>>>(
            h->res, h->rw, h->sel, h->begin, h->end, h->n, h->peer);
        CF_CUDA(cudaGetLastError());
        h->launches++;
        return launch_peer_barrier(h);      // every block has landed before the next fit pass
    }
    bisquare_kernel<false><<<dim3((cnt + 255) / 256, h->nseries), 256, 0, h->stream>>>(
        h->res, h->rw, h->sel, h->begin, h->end, h->n, PeerPtrs{});
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

// The two middle order statistics of an arbitrary device array (median_dbl,
// src/curvefit_statistics.f90:273-292), with the handle's select state as scratch:
// out = x(nmidp1) for odd n, half * (x(nmid) + x(nmidp1)) for even n.
__global__ void median_finish_kernel(const SelectState* st, int32_t n, double* out)
{
    const double lo = ordered_value(st->prefix[0]), hi = ordered_value(st->prefix[1]);
    *out = (n & 1) ? hi : __dmul_rn(0.5, __dadd_rn(lo, hi));
}

int select_median(cf_lowess* h, const double* d_x, int32_t n, double* d_out)
{
    select_reset_kernel<<<1, 256, 0, h->stream>>>(h->sel, h->hist, n);
    int blocks = (n + 512 * 8 - 1) / (512 * 8);
    if (blocks > 148 * 4) blocks = 148 * 4;
    for (int r = 0; r < kSelRounds; ++r) {
        hist_kernel<<<blocks, 512, 0, h->stream>>>(d_x, 0, n, n, r, h->sel, h->hist, 1);
        pick_kernel<false><<<1, 64, 0, h->stream>>>(h->sel, h->hist, r, PeerPtrs{}, 0ull);
    }
    median_finish_kernel<<<1, 1, 0, h->stream>>>(h->sel, n, d_out);
    CF_CUDA(cudaGetLastError());
    h->launches += 2 + 2 * kSelRounds;
    return CF_OK;
}

}  // namespace cf
