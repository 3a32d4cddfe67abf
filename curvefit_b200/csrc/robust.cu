// robust.cu -- the robustness step of lowess:
//   residuals + tie copy      src/curvefit_regression.f90:683-689, :695
//   median of |res|           :697-703  (full DLASRT sort in the reference;
//                             here an exact MSD radix SELECT of the two order
//                             statistics rw(m1), rw(m2) -- sorted values are
//                             algorithm independent, so the result is bit-equal)
//   bisquare weights          :704-713
//
// Passes over the data per robustness iteration (the reference sorts all of |res|):
//   A  residual_hist0_kernel  reads y, ys, leader, writes res  AND counts the top B0 bits
//                             of the key of |res| (shared-memory histogram per CTA); the
//                             last CTA to finish scans the counters and fixes the bucket
//                             of each of the two tracked ranks
//   B  compact_kernel         reads res, appends the keys of that bucket (1-3 % of n for
//                             B0 = 15) to a candidate list
//   C  cand_round_kernel      x `rounds`: W more bits per round, over the CANDIDATES only
//                             (kilobytes to megabytes); the last round forms cmad, c1, c9
//   D  bisquare_kernel        reads res, writes rw
// i.e. res is read twice and written once per iteration (was: 6 full histogram passes).
// All kernels work on the handle's shard [begin, end); across GPUs only the B0-bit
// histogram (an integer all-reduce done inside pass A's scan through peer memory) and
// the candidates (stored into every rank's list, peer_push_kernel) are exchanged, so every
// rank resolves the same two order statistics exactly.  Batched handles (nseries > 1):
// blockIdx.y is the series, every series has its own state.
#include "common.cuh"
#include "peer.cuh"

namespace cf {

namespace {

constexpr int kSelThreads = 1024;
constexpr unsigned kFullMask = 0xffffffffu;

enum { kSrcResidual = 0, kSrcAbs = 1, kSrcOrdered = 2 };

__device__ __forceinline__ unsigned long long abs_key(double v)
{
    return (unsigned long long)__double_as_longlong(fabs(v));
}

// Exclusive prefix of `v` over the CTA (1024 threads); `total` = sum over the CTA.
__device__ __forceinline__ unsigned long long block_exclusive(unsigned long long v, unsigned long long* sh33,
                                                              unsigned long long& total)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long t = __shfl_up_sync(kFullMask, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                         // sh33 may still be read from a previous call
    if (lane == 31) sh33[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        const int nw = blockDim.x >> 5;
        unsigned long long w = lane < nw ? sh33[lane] : 0ull, winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long t = __shfl_up_sync(kFullMask, winc, o);
            if (lane >= o) winc += t;
        }
        sh33[lane] = winc - w;
        if (lane == 31) sh33[32] = winc;
    }
    __syncthreads();
    total = sh33[32];
    return sh33[wid] + inc - v;
}

// Finds the bin of `cnt[0 .. bins)` that holds 0-based rank r (the counters sum to more than
// r) -- all threads of the CTA; the owner thread writes bin and the count below it.
__device__ __forceinline__ void block_pick(const unsigned int* cnt, int bins, unsigned long long r,
                                           unsigned long long* sh33, unsigned int* out_bin,
                                           unsigned long long* out_below)
{
    const int per = (bins + (int)blockDim.x - 1) / (int)blockDim.x;
    const int b0 = threadIdx.x * per;
    unsigned long long sum = 0;
    for (int k = 0; k < per; ++k)
        if (b0 + k < bins) sum += __ldcg(cnt + b0 + k);
    unsigned long long total;
    const unsigned long long exc = block_exclusive(sum, sh33, total);
    if (r >= exc && r < exc + sum) {
        unsigned long long cum = exc;
        for (int k = 0; k < per; ++k) {
            const unsigned long long c = __ldcg(cnt + b0 + k);
            if (r < cum + c) { *out_bin = (unsigned)(b0 + k); *out_below = cum; break; }
            cum += c;
        }
    }
    __syncthreads();
}

// last-block-done: true in every thread of the CTA that finished last (per series)
__device__ __forceinline__ bool last_block(unsigned int* ticket, unsigned int* sh_flag)
{
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(ticket, 1u);
        *sh_flag = (t == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    const bool last = *sh_flag != 0u;
    if (last) __threadfence();
    return last;
}

// Round 0, after the counters of all CTAs (and, PEER, of all ranks) are in hist0: the bucket
// of each tracked rank.  1-based m2 = n - m1 + 1 <= m1 = 1 + n/2  ->  0-based (n-1)/2 and
// n/2 (:699-700).  Leaves hist0 zeroed for the next iteration.
template <bool PEER>
__device__ __forceinline__ void pick0(SelectState* st, unsigned int* hist0, int32_t n, SelGeom g,
                                      const PeerPtrs& pp, unsigned long long epoch,
                                      unsigned long long* sh33)
{
    const int bins = 1 << g.b0;
    if (PEER) peer_hist_allreduce(pp, hist0, bins, epoch);
    __shared__ unsigned int s_bin[2];
    __shared__ unsigned long long s_below[2];
    const unsigned long long r0 = (unsigned long long)((n - 1) / 2), r1 = (unsigned long long)(n / 2);
    block_pick(hist0, bins, r0, sh33, &s_bin[0], &s_below[0]);
    block_pick(hist0, bins, r1, sh33, &s_bin[1], &s_below[1]);
    for (int k = threadIdx.x; k < bins; k += blockDim.x) hist0[k] = 0u;
    if (threadIdx.x == 0) {
        st->bucket[0] = s_bin[0]; st->bucket[1] = s_bin[1];
        st->prefix[0] = s_bin[0]; st->prefix[1] = s_bin[1];
        st->rank[0] = r0 - s_below[0]; st->rank[1] = r1 - s_below[1];
        st->ncand = 0u;
        st->ticket = 0u;
    }
}

// Round k >= 1 after the candidates' digits are counted: extend the prefixes; after the last
// round (LOWESS) form cmad, c1, c9 (:701-703).  Leaves hist zeroed.
__device__ __forceinline__ void pick_round(SelectState* st, unsigned int* hist, SelGeom g, int k, int finish,
                                           unsigned long long* sh33)
{
    const int width = sel_width(g, k), bins = 1 << width;
    const int hlen = 1 << g.w;
    __shared__ unsigned int s_bin[2];
    __shared__ unsigned long long s_below[2];
    const unsigned long long p0 = st->prefix[0], p1 = st->prefix[1];
    const unsigned long long r0 = st->rank[0], r1 = st->rank[1];
    const bool same = p0 == p1;
    block_pick(hist, bins, r0, sh33, &s_bin[0], &s_below[0]);
    block_pick(same ? hist : hist + hlen, bins, r1, sh33, &s_bin[1], &s_below[1]);
    for (int q = threadIdx.x; q < 2 * hlen; q += blockDim.x) hist[q] = 0u;
    if (threadIdx.x == 0) {
        const unsigned long long n0 = (p0 << width) | s_bin[0], n1 = (p1 << width) | s_bin[1];
        st->prefix[0] = n0; st->prefix[1] = n1;
        st->rank[0] = r0 - s_below[0]; st->rank[1] = r1 - s_below[1];
        st->ticket = 0u;
        if (k == g.rounds && finish) {
            const double lo = __longlong_as_double((long long)n0);               // rw(m2)
            const double hi = __longlong_as_double((long long)n1);               // rw(m1)
            const double cmad = __dmul_rn(3.0, __dadd_rn(hi, lo));               // :701
            st->cmad = cmad;
            st->c9 = __dmul_rn(0.999, cmad);                                     // :702
            st->c1 = __dmul_rn(0.001, cmad);                                     // :703
        }
    }
}

}  // namespace

// ---- residuals (last pass: no robustness step follows, :696) -----------------------------
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

// ---- pass A: residuals + round 0 of the select ---------------------------------------------
// SRC = kSrcResidual: the residual pass above, fused with the histogram of the top B0 key bits
// of |res|; kSrcAbs / kSrcOrdered: histogram only, of |src| / of the order-preserving image of
// the signed src (stats.cu).  defer: leave the counters in hist0 (a host collective sums them
// over the ranks; pick0_kernel follows).
template <int SRC, bool PEER>
__global__ void __launch_bounds__(kSelThreads, 1)
residual_hist0_kernel(const double* __restrict__ y, double* __restrict__ ys, double* __restrict__ res,
                      const int32_t* __restrict__ leader, int32_t begin, int32_t end, int32_t n,
                      int32_t ntotal, SelGeom g, SelectState* st, unsigned int* hist0, int defer,
                      const PeerPtrs pp, unsigned long long epoch)
{
    extern __shared__ unsigned int sh[];                 // 2^B0 counters
    __shared__ unsigned long long sh33[33];
    __shared__ unsigned int sh_flag;
    const int bins = 1 << g.b0;
    const int shift0 = g.top - g.b0;
    const size_t off = (size_t)blockIdx.y * (size_t)n;
    st += blockIdx.y;
    hist0 += (size_t)blockIdx.y * bins;
    for (int k = threadIdx.x; k < bins; k += blockDim.x) sh[k] = 0u;
    __syncthreads();
    // U independent elements per thread and trip: the loads of all U are in flight together
    // (one CTA of 1024 threads per SM -- the 128 KB histogram -- would otherwise be latency bound)
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)begin + blockIdx.x * blockDim.x + threadIdx.x; i0 < end; i0 += U * stride) {
        unsigned long long key[U];
        bool ok[U];
        if (SRC == kSrcResidual) {
            int32_t l[U];
            double v[U], yy[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + u * stride;
                ok[u] = i < end;
                l[u] = ok[u] ? max(leader[off + i], begin) : begin;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + u * stride;
                v[u] = ok[u] ? ys[off + l[u]] : 0.0;
                yy[u] = ok[u] ? y[off + i] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + u * stride;
                if (ok[u]) {
                    if (l[u] != (int32_t)i) ys[off + i] = v[u];                 // :685-688
                    const double r = __dsub_rn(yy[u], v[u]);                    // :695
                    res[off + i] = r;
                    key[u] = abs_key(r);                                        // :697
                }
            }
        } else {
            double v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + u * stride;
                ok[u] = i < end;
                v[u] = ok[u] ? res[off + i] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < U; ++u) key[u] = SRC == kSrcAbs ? abs_key(v[u]) : ordered_key(v[u]);
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (ok[u]) atomicAdd(&sh[(unsigned)(key[u] >> shift0)], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < bins; k += blockDim.x)
        if (sh[k]) atomicAdd(&hist0[k], sh[k]);
    if (!last_block(&st->ticket, &sh_flag)) return;
    if (defer) { if (threadIdx.x == 0) st->ticket = 0u; return; }
    pick0<PEER>(st, hist0, ntotal, g, pp, epoch, sh33);
}

// round 0 scan on its own (after a host collective summed hist0 over the ranks)
__global__ void __launch_bounds__(kSelThreads, 1)
pick0_kernel(SelectState* st, unsigned int* hist0, int32_t ntotal, SelGeom g)
{
    __shared__ unsigned long long sh33[33];
    st += blockIdx.x;
    hist0 += (size_t)blockIdx.x << g.b0;
    pick0<false>(st, hist0, ntotal, g, PeerPtrs{}, 0ull, sh33);
}

// ---- pass B: the keys of the bucket(s) holding the tracked ranks ----------------------------
// 1-3 % of the elements are candidates.  They are collected per CTA in shared memory (shared
// atomics) and appended to the list with ONE global atomic per flush: a global atomic per warp
// and trip on the single counter st->ncand serialised in L2 (n = 1e8: 1.47 ms for a pass whose
// 0.8 GB read takes 0.13 ms).
constexpr int kCompactThreads = 512, kCompactU = 4, kCompactBuf = 5120;
template <int SRC>
__global__ void __launch_bounds__(kCompactThreads)
compact_kernel(const double* __restrict__ src, int32_t begin, int32_t end, int32_t n, SelGeom g,
               SelectState* st, unsigned long long* __restrict__ cand, int64_t cand_stride,
               int64_t cand_base)
{
    __shared__ unsigned long long buf[kCompactBuf];
    __shared__ unsigned int cnt_sh, base_sh;
    src += (size_t)blockIdx.y * (size_t)n;
    st += blockIdx.y;
    cand += (size_t)blockIdx.y * (size_t)cand_stride + (size_t)cand_base;
    const int shift0 = g.top - g.b0;
    const unsigned b0 = st->bucket[0], b1 = st->bucket[1];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t cnt = (int64_t)end - begin;
    if (threadIdx.x == 0) cnt_sh = 0u;
    __syncthreads();
    auto flush = [&]() {                                     // all threads; cnt_sh is stable here
        const unsigned c = cnt_sh;
        if (threadIdx.x == 0) base_sh = c ? atomicAdd(&st->ncand, c) : 0u;
        __syncthreads();
        const unsigned base = base_sh;
        for (unsigned k = threadIdx.x; k < c; k += blockDim.x) cand[base + k] = buf[k];
        __syncthreads();
        if (threadIdx.x == 0) cnt_sh = 0u;
        __syncthreads();
    };
    // every thread of the CTA makes the same number of trips (the flush is a CTA-wide decision)
    const int64_t first = (int64_t)blockIdx.x * blockDim.x;
    for (int64_t k0 = first; k0 < cnt; k0 += kCompactU * stride) {
        double v[kCompactU];
        bool ok[kCompactU];
#pragma unroll
        for (int u = 0; u < kCompactU; ++u) {
            const int64_t k = k0 + threadIdx.x + u * stride;
            ok[u] = k < cnt;
            v[u] = ok[u] ? src[begin + k] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < kCompactU; ++u) {
            const unsigned long long key = SRC == kSrcOrdered ? ordered_key(v[u]) : abs_key(v[u]);
            const unsigned b = (unsigned)(key >> shift0);
            if (ok[u] && (b == b0 || b == b1)) buf[atomicAdd(&cnt_sh, 1u)] = key;
        }
        __syncthreads();
        if (cnt_sh > kCompactBuf - kCompactU * kCompactThreads) flush();
    }
    flush();
}

// ---- pass C: one more digit, over the candidates only ------------------------------------------
// Segments: one (the handle's own list, count = st->ncand) or, on peer-attached handles, one
// per rank (slot r of the list holds what rank r pushed, its count in PeerComm::cand_count).
__global__ void __launch_bounds__(kSelThreads, 1)
cand_round_kernel(const unsigned long long* __restrict__ cand, int64_t cand_stride, int nseg,
                  int64_t seg_cap, const unsigned int* __restrict__ seg_count, SelGeom g, int k,
                  SelectState* st, unsigned int* hist, int defer, int finish)
{
    __shared__ unsigned long long sh33[33];
    __shared__ unsigned int sh_flag;
    st += blockIdx.y;
    cand += (size_t)blockIdx.y * (size_t)cand_stride;
    const int hlen = 1 << g.w;
    hist += (size_t)blockIdx.y * 2 * hlen;
    const int width = sel_width(g, k), shift = sel_shift(g, k);
    const unsigned long long mask = (1ull << width) - 1ull;
    const unsigned long long p0 = st->prefix[0], p1 = st->prefix[1];
    const bool same = p0 == p1;
    const int hs = shift + width;
    // digits are counted in shared memory and only the non-zero counters go to the global
    // histogram: in the first round EVERY candidate matches the prefix, and one global atomic
    // each (3e6 at n = 1e8) made that round cost 0.15 ms
    static_assert(kSelWBatched <= kSelWSingle, "shared histogram is sized for the wider digit");
    __shared__ unsigned int sh_hist[2 << kSelWSingle];
    for (int q = threadIdx.x; q < 2 * hlen; q += blockDim.x) sh_hist[q] = 0u;
    __syncthreads();
    constexpr int U = 4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int s = 0; s < nseg; ++s) {
        const int64_t cnt = seg_count ? (int64_t)seg_count[s] : (int64_t)st->ncand;
        const unsigned long long* __restrict__ c = cand + (size_t)s * (size_t)seg_cap;
        for (int64_t q0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q0 < cnt; q0 += U * stride) {
            unsigned long long key[U];
#pragma unroll
            for (int u = 0; u < U; ++u) key[u] = q0 + u * stride < cnt ? c[q0 + u * stride] : 0ull;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (q0 + u * stride >= cnt) break;
                const unsigned long long hi = hs >= 64 ? 0ull : (key[u] >> hs);
                const unsigned d = (unsigned)((key[u] >> shift) & mask);
                if (hi == p0) atomicAdd(&sh_hist[d], 1u);
                if (!same && hi == p1) atomicAdd(&sh_hist[hlen + d], 1u);
            }
        }
    }
    __syncthreads();
    for (int q = threadIdx.x; q < 2 * hlen; q += blockDim.x)
        if (sh_hist[q]) atomicAdd(&hist[q], sh_hist[q]);
    if (!last_block(&st->ticket, &sh_flag)) return;
    if (defer) { if (threadIdx.x == 0) st->ticket = 0u; return; }
    pick_round(st, hist, g, k, finish, sh33);
}

__global__ void __launch_bounds__(kSelThreads, 1)
cand_pick_kernel(SelectState* st, unsigned int* hist, SelGeom g, int k, int finish)
{
    __shared__ unsigned long long sh33[33];
    st += blockIdx.x;
    hist += (size_t)blockIdx.x * 2 * ((size_t)1 << g.w);
    pick_round(st, hist, g, k, finish, sh33);
}

// ---- launchers -----------------------------------------------------------------------------------
static int sel_grid(int64_t cnt, int per_thread, int cap)
{
    int64_t b = (cnt + (int64_t)kSelThreads * per_thread - 1) / ((int64_t)kSelThreads * per_thread);
    if (b < 1) b = 1;
    if (b > cap) b = cap;
    return (int)b;
}

static int ensure_cand(cf_lowess* h)
{
    int64_t per = h->nranks > 1 ? (int64_t)h->nranks * h->block : (int64_t)h->n;
    if (per < 1) per = 1;
    const int64_t need = per * h->nseries;
    if (need > h->cap_cand) {
        peer_detach(h);                       // the list is one of the exported buffers
        cudaFree(h->cand);
        h->cand = nullptr;
        h->cap_cand = 0;
        CF_CUDA(cudaMalloc(&h->cand, sizeof(unsigned long long) * (size_t)need));
        h->cap_cand = need;
    }
    h->cand_stride = per;
    return CF_OK;
}

int select_prepare(cf_lowess* h) { return ensure_cand(h); }

static int set_hist0_smem(const void* fn)
{
    CF_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(sizeof(unsigned int) << kSelB0Single)));
    return CF_OK;
}

// pass A.  fused: with the residual pass (LOWESS, delta = 0); else over h->res as it stands.
int launch_hist0(cf_lowess* h, bool fused, bool defer)
{
    const SelGeom g = sel_geom(false, h->nseries > 1);
    const int64_t cnt = h->end > h->begin ? h->end - h->begin : 0;
    const int blocks = sel_grid(cnt, 8, h->nseries > 1 ? 16 : 148);
    const size_t smem = sizeof(unsigned int) << g.b0;
    const dim3 grid(blocks, h->nseries);
    const bool peer = h->peer_attached && !defer;
    if (!h->sel_attr) {
        CF_TRY(set_hist0_smem((const void*)residual_hist0_kernel<kSrcResidual, false>));
        CF_TRY(set_hist0_smem((const void*)residual_hist0_kernel<kSrcResidual, true>));
        CF_TRY(set_hist0_smem((const void*)residual_hist0_kernel<kSrcAbs, false>));
        CF_TRY(set_hist0_smem((const void*)residual_hist0_kernel<kSrcAbs, true>));
        h->sel_attr = true;
    }
    const unsigned long long epoch = peer ? ++h->epoch_hist : 0ull;
#define CF_A(SRC, PEER)                                                                              \
    residual_hist0_kernel<SRC, PEER><<<grid, kSelThreads, smem, h->stream>>>(                        \
        h->y, h->ys, h->res, h->win.leader, h->begin, h->end, h->n, h->n, g, h->sel, h->hist0,       \
        defer ? 1 : 0, peer ? h->peer : PeerPtrs{}, epoch)
    if (fused) { if (peer) CF_A(kSrcResidual, true); else CF_A(kSrcResidual, false); }
    else       { if (peer) CF_A(kSrcAbs, true); else CF_A(kSrcAbs, false); }
#undef CF_A
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_pick0(cf_lowess* h)
{
    const SelGeom g = sel_geom(false, h->nseries > 1);
    pick0_kernel<<<h->nseries, kSelThreads, 0, h->stream>>>(h->sel, h->hist0, h->n, g);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

// pass B (+ on peer-attached handles the exchange of the candidates)
int launch_compact(cf_lowess* h)
{
    const SelGeom g = sel_geom(false, h->nseries > 1);
    CF_TRY(ensure_cand(h));
    const int64_t cnt = h->end > h->begin ? h->end - h->begin : 0;
    if (cnt > 0) {
        int64_t b = (cnt + kCompactThreads * 8 - 1) / (kCompactThreads * 8);
        const int cap = h->nseries > 1 ? 16 : 148 * 4;
        if (b > cap) b = cap;
        const int64_t base = h->nranks > 1 ? (int64_t)h->rank * h->block : 0;
        compact_kernel<kSrcAbs><<<dim3((int)b, h->nseries), kCompactThreads, 0, h->stream>>>(
            h->res, h->begin, h->end, h->n, g, h->sel, h->cand, h->cand_stride, base);
        CF_CUDA(cudaGetLastError());
        h->launches++;
    }
    if (h->peer_attached) return launch_peer_push(h);
    return CF_OK;
}

// pass C, round k = 1 .. rounds
int launch_cand_round(cf_lowess* h, int k, bool defer)
{
    const SelGeom g = sel_geom(false, h->nseries > 1);
    const bool peer = h->peer_attached && !defer;
    // the candidate count lives on the device; a grid sized for ~2 % of the shard is plenty and
    // harmless when there are fewer
    const int64_t guess = ((int64_t)h->n * (h->nseries > 1 ? 2 : 1)) / 32 + 1;
    const int blocks = sel_grid(guess, 4, h->nseries > 1 ? 4 : 64);     // few CTAs: fewer counters to flush
    const int nseg = h->nranks > 1 ? (peer ? h->nranks : 1) : 1;
    // own slot of a sharded handle starts at rank * block
    const unsigned long long* cand = h->cand + ((h->nranks > 1 && !peer) ? (size_t)h->rank * h->block : 0);
    cand_round_kernel<<<dim3(blocks, h->nseries), kSelThreads, 0, h->stream>>>(
        cand, h->cand_stride, nseg, (int64_t)h->block, peer ? h->comm->cand_count : nullptr, g, k,
        h->sel, h->hist, defer ? 1 : 0, 1);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_cand_pick(cf_lowess* h, int k)
{
    const SelGeom g = sel_geom(false, h->nseries > 1);
    cand_pick_kernel<<<h->nseries, kSelThreads, 0, h->stream>>>(h->sel, h->hist, g, k, 1);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int select_rounds(const cf_lowess* h) { return sel_geom(false, h->nseries > 1).rounds; }

// the whole select after a fit pass (single GPU, or peer-attached: the exchanges are inside)
int launch_select(cf_lowess* h, bool fused_residuals)
{
    CF_TRY(launch_hist0(h, fused_residuals, false));
    CF_TRY(launch_compact(h));
    const int rounds = select_rounds(h);
    for (int k = 1; k <= rounds; ++k) CF_TRY(launch_cand_round(h, k, false));
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
// src/curvefit_statistics.f90:273-292) on `stream`, with caller-provided scratch:
// out = x(nmidp1) for odd n, half * (x(nmid) + x(nmidp1)) for even n.
__global__ void median_finish_kernel(const SelectState* st, int32_t n, double* out)
{
    const double lo = ordered_value(st->prefix[0]), hi = ordered_value(st->prefix[1]);
    *out = (n & 1) ? hi : __dmul_rn(0.5, __dadd_rn(lo, hi));
}

size_t select_scratch_bytes(int32_t n)
{
    // state | hist0 | hist | candidates
    return 256 + (sizeof(unsigned int) << kSelB0Single) + 2 * (sizeof(unsigned int) << kSelWSingle)
           + sizeof(unsigned long long) * (size_t)(n > 0 ? n : 1);
}

int select_median(const double* d_x, int32_t n, double* d_out, void* scratch, cudaStream_t s, int* launches)
{
    const SelGeom g = sel_geom(true, false);
    unsigned char* p = static_cast<unsigned char*>(scratch);
    SelectState* st = reinterpret_cast<SelectState*>(p);
    unsigned int* hist0 = reinterpret_cast<unsigned int*>(p + 256);
    unsigned int* hist = hist0 + (1 << kSelB0Single);
    unsigned long long* cand = reinterpret_cast<unsigned long long*>(hist + 2 * (1 << kSelWSingle));
    CF_CUDA(cudaMemsetAsync(scratch, 0, 256 + (sizeof(unsigned int) << kSelB0Single)
                                            + 2 * (sizeof(unsigned int) << kSelWSingle), s));
    CF_TRY(set_hist0_smem((const void*)residual_hist0_kernel<kSrcOrdered, false>));
    const int blocks = sel_grid(n, 8, 148);
    residual_hist0_kernel<kSrcOrdered, false><<<blocks, kSelThreads, sizeof(unsigned int) << g.b0, s>>>(
        nullptr, nullptr, const_cast<double*>(d_x), nullptr, 0, n, n, n, g, st, hist0, 0, PeerPtrs{}, 0ull);
    int64_t b = ((int64_t)n + 512 * 8 - 1) / (512 * 8);
    if (b > 148 * 4) b = 148 * 4;
    compact_kernel<kSrcOrdered><<<(int)b, 512, 0, s>>>(d_x, 0, n, n, g, st, cand, n, 0);
    const int cblocks = sel_grid((int64_t)n / 32 + 1, 4, 148);
    for (int k = 1; k <= g.rounds; ++k)
        cand_round_kernel<<<cblocks, kSelThreads, 0, s>>>(cand, n, 1, n, nullptr, g, k, st, hist, 0, 0);
    median_finish_kernel<<<1, 1, 0, s>>>(st, n, d_out);
    CF_CUDA(cudaGetLastError());
    if (launches) *launches += 3 + g.rounds;
    return CF_OK;
}

}  // namespace cf
