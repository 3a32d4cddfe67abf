// sort.cu -- ls_init's ascending sort of x with y co-permuted (SURVEY 8a rows
// a2/a3; src/curvefit_regression.f90:793-798, which calls LINALG's recursive
// first-pivot quicksort, linalg_sorting.f90:277-354).
//
// Here: a stable LSD radix sort on the device, 8 passes of 8 bits over an
// order-preserving uint64 image of the doubles, carrying the uint32 source
// index.  Sorted x VALUES are a function of the multiset of x only, so they are
// bit-identical to any correct sort; the order of y inside a group of equal x
// is by source index here (stable), quicksort-specific in the reference.
//
// Per pass: (1) per-tile digit histogram, (2) exclusive scan of the
// digit-major [256][tiles] table, (3) stable scatter with warp-level
// match-any ranking.  HBM-bound: reads 2 x 12 B, writes 12 B per element.
#include "common.cuh"

namespace cf {

constexpr int kSortThreads = 256;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortKpt = 16;                        // keys per thread
constexpr int kSortTile = kSortThreads * kSortKpt;  // 4096 keys per CTA
constexpr int kScanThreads = 1024;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ unsigned long long key_of(double v)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double val_of(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

__global__ void sort_prepare_kernel(const double* __restrict__ x, int32_t n,
                                    unsigned long long* __restrict__ keys,
                                    uint32_t* __restrict__ idx)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { keys[i] = key_of(x[i]); idx[i] = (uint32_t)i; }
}

__global__ void __launch_bounds__(kSortThreads)
sort_hist_kernel(const unsigned long long* __restrict__ keys, int32_t n, int shift,
                 uint32_t* __restrict__ table, int ntiles)
{
    __shared__ unsigned int sh[256];
    sh[threadIdx.x] = 0u;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * kSortTile;
#pragma unroll 4
    for (int k = 0; k < kSortKpt; ++k) {
        const int64_t i = base + k * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&sh[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    table[(size_t)threadIdx.x * ntiles + blockIdx.x] = sh[threadIdx.x];
}

// ---- exclusive scan of uint32 (two levels: <= kScanTile^2 entries) ------------------
__global__ void __launch_bounds__(kScanThreads)
scan_tile_kernel(uint32_t* __restrict__ data, int64_t len, uint32_t* __restrict__ sums)
{
    __shared__ uint32_t wsum[32];
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
    uint32_t v[kScanItems], tot = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        v[k] = (base + k < len) ? data[base + k] : 0u;
        tot += v[k];
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint32_t inc = tot;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) wsum[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        uint32_t w = wsum[lane], wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        wsum[lane] = wi - w;
        if (lane == 31 && sums) sums[blockIdx.x] = wi;
    }
    __syncthreads();
    uint32_t run = wsum[wid] + inc - tot;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k) {
        if (base + k < len) data[base + k] = run;
        run += v[k];
    }
}

__global__ void __launch_bounds__(kScanThreads)
scan_add_kernel(uint32_t* __restrict__ data, int64_t len, const uint32_t* __restrict__ sums)
{
    const uint32_t add = sums[blockIdx.x];
    const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; ++k)
        if (base + k < len) data[base + k] += add;
}

// ---- stable scatter ----------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads)
sort_scatter_kernel(const unsigned long long* __restrict__ keys_in,
                    const uint32_t* __restrict__ idx_in, int32_t n, int shift,
                    const uint32_t* __restrict__ table, int ntiles,
                    unsigned long long* __restrict__ keys_out, uint32_t* __restrict__ idx_out)
{
    __shared__ unsigned int cnt[kSortWarps][256];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < kSortWarps * 256; k += kSortThreads) (&cnt[0][0])[k] = 0u;
    __syncthreads();

    // warp `wid` owns a contiguous run of 32*kSortKpt keys, visited round by round
    const int64_t wbase = (int64_t)blockIdx.x * kSortTile + (int64_t)wid * (32 * kSortKpt);
    unsigned long long key[kSortKpt];
    uint32_t val[kSortKpt];
    unsigned int rank[kSortKpt];
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        const bool ok = i < n;
        key[r] = ok ? keys_in[i] : ~0ull;
        val[r] = ok ? idx_in[i] : 0u;
        const unsigned d = ok ? ((unsigned)(key[r] >> shift) & 255u) : 256u;
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const int first = __ffs(peers) - 1;
        unsigned old = 0;
        if (ok && lane == first) {
            old = cnt[wid][d];
            cnt[wid][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, first);
        rank[r] = old + __popc(peers & lt);
        __syncwarp();
    }
    __syncthreads();
    {   // per digit: global tile offset + exclusive scan over the CTA's warps
        const int d = threadIdx.x;
        unsigned run = table[(size_t)d * ntiles + blockIdx.x];
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const unsigned c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        if (i < n) {
            const unsigned d = (unsigned)(key[r] >> shift) & 255u;
            const unsigned pos = cnt[wid][d] + rank[r];
            keys_out[pos] = key[r];
            idx_out[pos] = val[r];
        }
    }
}

__global__ void sort_finish_kernel(const unsigned long long* __restrict__ keys,
                                   const uint32_t* __restrict__ idx,
                                   const double* __restrict__ y_in, int32_t n,
                                   double* __restrict__ x_out, double* __restrict__ y_out,
                                   int32_t* __restrict__ perm)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = idx[i];
    x_out[i] = val_of(keys[i]);
    if (y_out) y_out[i] = y_in[s];               // m_y = y(indices), :798
    if (perm) perm[i] = (int32_t)s + 1;          // 1-based like the tracking array
}

int sort_xy_device(const double* d_x_in, const double* d_y_in, int32_t n, double* d_x_out,
                   double* d_y_out, int32_t* d_perm_out, cudaStream_t s, int* launches)
{
    if (n <= 0) return CF_OK;
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    const int64_t tlen = (int64_t)256 * ntiles;
    const int nscan = (int)((tlen + kScanTile - 1) / kScanTile);
    if (nscan > kScanTile) { set_error("sort: n too large for the two-level scan"); return CF_INVALID_INPUT_ERROR; }

    unsigned long long *k0 = nullptr, *k1 = nullptr;
    uint32_t *v0 = nullptr, *v1 = nullptr, *table = nullptr, *sums = nullptr;
    cudaError_t e = cudaSuccess;
    auto release = [&]() {
        cudaFree(k0); cudaFree(k1); cudaFree(v0); cudaFree(v1); cudaFree(table); cudaFree(sums);
    };
    if ((e = cudaMalloc(&k0, sizeof(unsigned long long) * (size_t)n)) != cudaSuccess ||
        (e = cudaMalloc(&k1, sizeof(unsigned long long) * (size_t)n)) != cudaSuccess ||
        (e = cudaMalloc(&v0, sizeof(uint32_t) * (size_t)n)) != cudaSuccess ||
        (e = cudaMalloc(&v1, sizeof(uint32_t) * (size_t)n)) != cudaSuccess ||
        (e = cudaMalloc(&table, sizeof(uint32_t) * (size_t)tlen)) != cudaSuccess ||
        (e = cudaMalloc(&sums, sizeof(uint32_t) * (size_t)kScanTile)) != cudaSuccess) {
        release();
        return cuda_fail(e, "cudaMalloc(sort scratch)", __FILE__, __LINE__);
    }
    int nl = 0;
    const int pb = (n + 255) / 256;
    sort_prepare_kernel<<<pb, 256, 0, s>>>(d_x_in, n, k0, v0); ++nl;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        sort_hist_kernel<<<ntiles, kSortThreads, 0, s>>>(k0, n, shift, table, ntiles); ++nl;
        scan_tile_kernel<<<nscan, kScanThreads, 0, s>>>(table, tlen, sums); ++nl;
        if (nscan > 1) {
            scan_tile_kernel<<<1, kScanThreads, 0, s>>>(sums, nscan, nullptr); ++nl;
            scan_add_kernel<<<nscan, kScanThreads, 0, s>>>(table, tlen, sums); ++nl;
        }
        sort_scatter_kernel<<<ntiles, kSortThreads, 0, s>>>(k0, v0, n, shift, table, ntiles, k1, v1);
        ++nl;
        unsigned long long* tk = k0; k0 = k1; k1 = tk;
        uint32_t* tv = v0; v0 = v1; v1 = tv;
    }
    sort_finish_kernel<<<pb, 256, 0, s>>>(k0, v0, d_y_in, n, d_x_out, d_y_out, d_perm_out); ++nl;
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    release();
    if (launches) *launches += nl;
    if (e != cudaSuccess) return cuda_fail(e, "radix sort", __FILE__, __LINE__);
    return CF_OK;
}

}  // namespace cf
