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
constexpr int kSortMinBlocks = 3;                   // 3 x (4096 x 16 B + 10 KB) of shared memory per SM
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
// V = payload carried with the key: the uint32 source index (cf_sort_xy, which returns the
// permutation) or the 8 bytes of y itself (initialize: no gather afterwards).
template <typename V>
__global__ void __launch_bounds__(kSortThreads, kSortMinBlocks)
sort_scatter_kernel(const unsigned long long* __restrict__ keys_in,
                    const V* __restrict__ idx_in, int32_t n, int shift,
                    const uint32_t* __restrict__ table, int ntiles,
                    unsigned long long* __restrict__ keys_out, V* __restrict__ idx_out)
{
    extern __shared__ __align__(16) unsigned char sort_smem[];
    unsigned long long* skey = reinterpret_cast<unsigned long long*>(sort_smem);          // [kSortTile]
    V* sval = reinterpret_cast<V*>(sort_smem + sizeof(unsigned long long) * kSortTile);
    __shared__ unsigned int cnt[kSortWarps][256];
    __shared__ unsigned int dbase[256], gofs[256], wtot[kSortWarps];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < kSortWarps * 256; k += kSortThreads) (&cnt[0][0])[k] = 0u;
    __syncthreads();

    // warp `wid` owns a contiguous run of 32*kSortKpt keys, visited round by round
    const int64_t wbase = (int64_t)blockIdx.x * kSortTile + (int64_t)wid * (32 * kSortKpt);
    unsigned long long key[kSortKpt];
    unsigned int rank[kSortKpt], peers[kSortKpt];
    const unsigned lt = (1u << lane) - 1u;
    // all loads and all lane matches first (independent, pipelined) ...
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        const bool ok = i < n;
        key[r] = ok ? keys_in[i] : ~0ull;
    }
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const bool ok = wbase + r * 32 + lane < n;
        const unsigned d = ok ? ((unsigned)(key[r] >> shift) & 255u) : 256u;
        peers[r] = __match_any_sync(0xffffffffu, d);
    }
    // ... then the warp's digit counters, round by round (stable: round-major, lane-minor)
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const bool ok = wbase + r * 32 + lane < n;
        const unsigned d = (unsigned)(key[r] >> shift) & 255u;
        const int first = __ffs(peers[r]) - 1;
        unsigned old = 0;
        if (ok && lane == first) {
            old = cnt[wid][d];
            cnt[wid][d] = old + __popc(peers[r]);
        }
        old = __shfl_sync(0xffffffffu, old, first);
        rank[r] = old + __popc(peers[r] & lt);
        __syncwarp();
    }
    __syncthreads();
    // Per digit (thread d): exclusive scan of the warps' counts, then a CTA-wide
    // exclusive scan of the digit totals = where the digit's run starts inside the
    // tile once it is ordered by digit; gofs[d] + (slot in the ordered tile) = the
    // key's final position.
    {
        const int d = threadIdx.x;
        unsigned run = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) {
            const unsigned c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
        unsigned inc = run;                                  // inclusive scan over the 256 digits
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) wtot[wid] = inc;
        __syncthreads();
        unsigned before = 0;
#pragma unroll
        for (int w = 0; w < kSortWarps; ++w) before += (w < wid) ? wtot[w] : 0u;
        const unsigned lbase = before + inc - run;
        dbase[d] = lbase;
        gofs[d] = table[(size_t)d * ntiles + blockIdx.x] - lbase;
    }
    __syncthreads();
    // order the tile by digit in shared memory (stable) ...
#pragma unroll
    for (int r = 0; r < kSortKpt; ++r) {
        const int64_t i = wbase + r * 32 + lane;
        if (i < n) {
            const unsigned d = (unsigned)(key[r] >> shift) & 255u;
            const unsigned pos = dbase[d] + cnt[wid][d] + rank[r];
            skey[pos] = key[r];
            sval[pos] = idx_in[i];          // the payload goes straight from global to its slot
        }
    }
    __syncthreads();
    // ... and write it out: consecutive threads hold consecutive slots of a digit's run,
    // so every run leaves as one contiguous (coalesced) piece
    const int64_t left = (int64_t)n - (int64_t)blockIdx.x * kSortTile;
    const int tile_n = left < kSortTile ? (int)left : kSortTile;
#pragma unroll 4
    for (int t = threadIdx.x; t < tile_n; t += kSortThreads) {
        const unsigned long long k = skey[t];
        const unsigned pos = gofs[(unsigned)(k >> shift) & 255u] + (unsigned)t;
        keys_out[pos] = k;
        idx_out[pos] = sval[t];
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

__global__ void sort_keys_inplace_kernel(double* __restrict__ x, int32_t n)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) reinterpret_cast<unsigned long long*>(x)[i] = key_of(x[i]);
}

__global__ void sort_unkey_kernel(const unsigned long long* __restrict__ keys,
                                  const unsigned long long* __restrict__ ybits, int32_t n,
                                  double* __restrict__ x_out, double* __restrict__ y_out)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    x_out[i] = val_of(keys[i]);
    y_out[i] = __longlong_as_double((long long)ybits[i]);       // m_y = y(indices), :798
}

// 8 LSD passes of 8 bits; the result is back in (k0, v0)
template <typename V>
static int sort_passes(unsigned long long* k0, V* v0, unsigned long long* k1, V* v1, int32_t n,
                       uint32_t* table, uint32_t* sums, cudaStream_t s, int* nl)
{
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    const int64_t tlen = (int64_t)256 * ntiles;
    const int nscan = (int)((tlen + kScanTile - 1) / kScanTile);
    constexpr size_t smem = (sizeof(unsigned long long) + sizeof(V)) * kSortTile;
    CF_CUDA(cudaFuncSetAttribute(sort_scatter_kernel<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        sort_hist_kernel<<<ntiles, kSortThreads, 0, s>>>(k0, n, shift, table, ntiles); ++*nl;
        scan_tile_kernel<<<nscan, kScanThreads, 0, s>>>(table, tlen, sums); ++*nl;
        if (nscan > 1) {
            scan_tile_kernel<<<1, kScanThreads, 0, s>>>(sums, nscan, nullptr); ++*nl;
            scan_add_kernel<<<nscan, kScanThreads, 0, s>>>(table, tlen, sums); ++*nl;
        }
        sort_scatter_kernel<V><<<ntiles, kSortThreads, smem, s>>>(k0, v0, n, shift, table, ntiles, k1, v1);
        ++*nl;
        unsigned long long* tk = k0; k0 = k1; k1 = tk;
        V* tv = v0; v0 = v1; v1 = tv;
    }
    CF_CUDA(cudaGetLastError());
    return CF_OK;
}

size_t sort_aux_bytes(int32_t n)
{
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    return sizeof(uint32_t) * ((size_t)256 * ntiles + kScanTile);
}

// `pairs`: 24 n bytes for the two (key, index) buffers, `aux`: sort_aux_bytes(n) for the digit
// table; either may be null, then the scratch is allocated and freed here.
int sort_xy_device(const double* d_x_in, const double* d_y_in, int32_t n, double* d_x_out,
                   double* d_y_out, int32_t* d_perm_out, cudaStream_t s, int* launches,
                   void* pairs, void* aux)
{
    if (n <= 0) return CF_OK;
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    const int64_t tlen = (int64_t)256 * ntiles;
    const int nscan = (int)((tlen + kScanTile - 1) / kScanTile);
    if (nscan > kScanTile) { set_error("sort: n too large for the two-level scan"); return CF_INVALID_INPUT_ERROR; }

    unsigned long long *k0 = nullptr, *k1 = nullptr;
    uint32_t *v0 = nullptr, *v1 = nullptr, *table = nullptr, *sums = nullptr;
    void *own_pairs = nullptr, *own_aux = nullptr;
    cudaError_t e = cudaSuccess;
    auto release = [&]() { cudaFree(own_pairs); cudaFree(own_aux); };
    if (!pairs) {
        if ((e = cudaMalloc(&own_pairs, (size_t)24 * (size_t)n)) != cudaSuccess)
            return cuda_fail(e, "cudaMalloc(sort scratch)", __FILE__, __LINE__);
        pairs = own_pairs;
    }
    if (!aux) {
        if ((e = cudaMalloc(&own_aux, sort_aux_bytes(n))) != cudaSuccess) {
            release();
            return cuda_fail(e, "cudaMalloc(sort scratch)", __FILE__, __LINE__);
        }
        aux = own_aux;
    }
    k0 = static_cast<unsigned long long*>(pairs);
    k1 = k0 + n;
    v0 = reinterpret_cast<uint32_t*>(k1 + n);
    v1 = v0 + n;
    table = static_cast<uint32_t*>(aux);
    sums = table + tlen;
    int nl = 0;
    const int pb = (n + 255) / 256;
    sort_prepare_kernel<<<pb, 256, 0, s>>>(d_x_in, n, k0, v0); ++nl;
    {
        const int rc = sort_passes<uint32_t>(k0, v0, k1, v1, n, table, sums, s, &nl);
        if (rc != CF_OK) { release(); return rc; }
    }
    sort_finish_kernel<<<pb, 256, 0, s>>>(k0, v0, d_y_in, n, d_x_out, d_y_out, d_perm_out); ++nl;
    e = cudaGetLastError();
    if (e == cudaSuccess && (own_pairs || own_aux)) e = cudaStreamSynchronize(s);   // before the scratch goes
    release();
    if (launches) *launches += nl;
    if (e != cudaSuccess) return cuda_fail(e, "radix sort", __FILE__, __LINE__);
    return CF_OK;
}

// initialize(): x_io / y_io hold the unsorted series and are used as one half of the
// ping-pong (clobbered); `scratch` = 16 n bytes for the other half; y travels with its key
// as the payload, so nothing is gathered at the end.  Asynchronous on s.
int sort_xy_inplace(double* x_io, double* y_io, int32_t n, double* x_out, double* y_out,
                    void* scratch, void* aux, cudaStream_t s, int* launches)
{
    if (n <= 0) return CF_OK;
    const int ntiles = (n + kSortTile - 1) / kSortTile;
    const int64_t tlen = (int64_t)256 * ntiles;
    if ((tlen + kScanTile - 1) / kScanTile > kScanTile) { set_error("sort: n too large for the two-level scan"); return CF_INVALID_INPUT_ERROR; }
    unsigned long long* k0 = reinterpret_cast<unsigned long long*>(x_io);
    unsigned long long* v0 = reinterpret_cast<unsigned long long*>(y_io);
    unsigned long long* k1 = static_cast<unsigned long long*>(scratch);
    unsigned long long* v1 = k1 + n;
    uint32_t* table = static_cast<uint32_t*>(aux);
    uint32_t* sums = table + tlen;
    int nl = 0;
    const int pb = (n + 255) / 256;
    sort_keys_inplace_kernel<<<pb, 256, 0, s>>>(x_io, n); ++nl;
    CF_TRY(sort_passes<unsigned long long>(k0, v0, k1, v1, n, table, sums, s, &nl));
    sort_unkey_kernel<<<pb, 256, 0, s>>>(k0, v0, n, x_out, y_out); ++nl;
    CF_CUDA(cudaGetLastError());
    if (launches) *launches += nl;
    return CF_OK;
}

}  // namespace cf
