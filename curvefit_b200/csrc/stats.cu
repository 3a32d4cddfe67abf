// stats.cu -- the routines next to the LOWESS path:
//   cf_moving_average   moving_average_1, src/curvefit_regression.f90:427-478
//   cf_median           median_dbl,       src/curvefit_statistics.f90:258-295
//   cf_sort             sort(x, .true.),  LINALG linalg_sorting.f90:12-35 -> DLASRT('I')
//
// moving_average.  The reference keeps the last npts samples in a zero-initialised
// shift buffer (newest first) and replaces x(i) by mean(buffer) (:467-477), where
// mean is the running update z = b(1); z = z + (b(k) - z)/k, k = 2..npts
// (curvefit_statistics.f90:239-242).  The first npts-1 outputs therefore average in
// zeros, and the rounding of every output is that of this particular recurrence --
// not of a windowed sum.  The kernel evaluates the same recurrence, one thread per
// output, with the reference's operation order and IEEE division, so the result is
// BIT-IDENTICAL to the reference's.  A tile of NT outputs walks the NT + C - 1 input
// samples of each chunk of C steps through shared memory (conflict-free: at step k
// thread t reads sample i_t - k + 1).  Work is n * npts updates like the reference's;
// cost per update = one FP64 subtract, divide, add.
#include "common.cuh"

#include <mutex>

namespace cf {

template <int NT, int C>
__global__ void __launch_bounds__(NT)
moving_average_kernel(const double* __restrict__ x, double* __restrict__ out, int32_t n, int32_t npts)
{
    __shared__ double buf[NT + C];
    const int tid = threadIdx.x;
    const int64_t tile0 = (int64_t)blockIdx.x * NT;
    const int64_t i = tile0 + tid;
    double z = i < n ? x[i] : 0.0;                              // buffer(1) = x(i), z = b(1)
    // chunk s covers steps k = s*C + 2 .. s*C + C + 1, i.e. samples
    // j = i - k + 1 in [tile0 - (s+1)*C, tile0 + NT - 2 - s*C]
    for (int32_t k0 = 2; k0 <= npts; k0 += C) {
        const int64_t base = tile0 - (k0 - 2) - C;              // sample held in buf[0]
        __syncthreads();
        for (int m = tid; m < NT + C - 1; m += NT) {
            const int64_t j = base + m;
            buf[m] = (j >= 0 && j < n) ? x[j] : 0.0;            // buffer = zero before the signal starts
        }
        __syncthreads();
        const int steps = min(C, npts - k0 + 1);
        // (tile0 + NT <= the first sample of the chunk for every thread: only zeros remain)
#pragma unroll 4
        for (int kk = 0; kk < steps; ++kk) {
            const double b = buf[tid + C - 1 - kk];
            const double k = (double)(k0 + kk);
            z = __dadd_rn(z, __ddiv_rn(__dsub_rn(b, z), k));    // z = z + (x(i) - z) / i
        }
    }
    if (i < n) out[i] = z;
}

int moving_average_device(const double* d_x, double* d_out, int32_t n, int32_t npts, cudaStream_t s)
{
    constexpr int NT = 128, C = 256;
    const int blocks = (n + NT - 1) / NT;
    moving_average_kernel<NT, C><<<blocks, NT, 0, s>>>(d_x, d_out, n, npts);
    CF_CUDA(cudaGetLastError());
    return CF_OK;
}

}  // namespace cf

using namespace cf;

extern "C" {

int32_t cf_moving_average(double* x, int32_t n, int32_t npts)
{
    if (n < 0 || (n > 0 && !x)) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (npts < 2 || npts > n) {                                  // :452-457
        set_error("The averaging size window must be at least 2, and less than the size of the data set.");
        return CF_INVALID_INPUT_ERROR;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available; libcurvefit_cuda has no CPU fallback");
        return CF_DEVICE_ERROR;
    }
    double *dx = nullptr, *dout = nullptr;
    const size_t bytes = sizeof(double) * (size_t)n;
    cudaError_t e = cudaMalloc(&dx, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&dout, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(dx, x, bytes, cudaMemcpyHostToDevice);
    int rc = e == cudaSuccess ? moving_average_device(dx, dout, n, npts, nullptr)
                              : cuda_fail(e, "cf_moving_average upload", __FILE__, __LINE__);
    if (rc == CF_OK) {
        e = cudaMemcpy(x, dout, bytes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = cuda_fail(e, "cf_moving_average download", __FILE__, __LINE__);
    }
    cudaFree(dx); cudaFree(dout);
    return rc;
}

int32_t cf_moving_average_device(const void* d_x, void* d_out, int32_t n, int32_t npts, void* stream)
{
    if (n < 0 || (n > 0 && (!d_x || !d_out)) || d_x == d_out) {
        set_error("Input array sizes must match (in and out may not alias).");
        return CF_ARRAY_SIZE_ERROR;
    }
    if (npts < 2 || npts > n) {
        set_error("The averaging size window must be at least 2, and less than the size of the data set.");
        return CF_INVALID_INPUT_ERROR;
    }
    return moving_average_device(static_cast<const double*>(d_x), static_cast<double*>(d_out), n, npts,
                                 (cudaStream_t)stream);
}

int32_t cf_sort(const double* x, int32_t n, double* x_sorted)
{
    if (n < 0 || (n > 0 && (!x || !x_sorted))) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (n == 0) return CF_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available; libcurvefit_cuda has no CPU fallback");
        return CF_DEVICE_ERROR;
    }
    double *dx = nullptr, *ox = nullptr;
    const size_t bytes = sizeof(double) * (size_t)n;
    cudaError_t e = cudaMalloc(&dx, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&ox, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(dx, x, bytes, cudaMemcpyHostToDevice);
    int rc = e == cudaSuccess ? sort_xy_device(dx, nullptr, n, ox, nullptr, nullptr, nullptr, nullptr)
                              : cuda_fail(e, "cf_sort upload", __FILE__, __LINE__);
    if (rc == CF_OK) {
        e = cudaMemcpy(x_sorted, ox, bytes, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = cuda_fail(e, "cf_sort download", __FILE__, __LINE__);
    }
    cudaFree(dx); cudaFree(ox);
    return rc;
}

int32_t cf_median(const double* x, int32_t n, double* z)
{
    if (n < 0 || !z || (n > 0 && !x)) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (n == 0) { *z = 0.0; return CF_OK; }                      // :280-283
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available; libcurvefit_cuda has no CPU fallback");
        return CF_DEVICE_ERROR;
    }
    // one allocation: data | result | select scratch
    const size_t bytes = sizeof(double) * (size_t)n;
    unsigned char* d = nullptr;
    cudaError_t e = cudaMalloc(&d, bytes + 256 + select_scratch_bytes(n));
    if (e != cudaSuccess) return cuda_fail(e, "cf_median alloc", __FILE__, __LINE__);
    double* dx = reinterpret_cast<double*>(d);
    double* dz = reinterpret_cast<double*>(d + bytes);
    e = cudaMemcpyAsync(dx, x, bytes, cudaMemcpyHostToDevice, cudaStreamPerThread);
    int rc = e == cudaSuccess ? select_median(dx, n, dz, d + bytes + 256, cudaStreamPerThread, nullptr)
                              : cuda_fail(e, "cf_median upload", __FILE__, __LINE__);
    if (rc == CF_OK) {
        e = cudaMemcpyAsync(z, dz, sizeof(double), cudaMemcpyDeviceToHost, cudaStreamPerThread);
        if (e == cudaSuccess) e = cudaStreamSynchronize(cudaStreamPerThread);
        if (e != cudaSuccess) rc = cuda_fail(e, "cf_median", __FILE__, __LINE__);
    }
    cudaFree(d);
    return rc;
}

// Asynchronous on `stream` (a cudaStream_t; NULL = the legacy default stream, like
// cf_moving_average_device): the scratch is allocated and freed in stream order.
int32_t cf_median_device(const void* d_x, int32_t n, void* d_z, void* stream)
{
    if (n < 1 || !d_x || !d_z) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    cudaStream_t s = (cudaStream_t)stream;
    void* scratch = nullptr;
    CF_CUDA(cudaMallocAsync(&scratch, select_scratch_bytes(n), s));
    const int rc = select_median(static_cast<const double*>(d_x), n, static_cast<double*>(d_z), scratch, s, nullptr);
    cudaFreeAsync(scratch, s);
    return rc;
}

}  // extern "C"

// ---- cf_median_device_noalloc ---------------------------------------------------------------
// The select's scratch (state | two histograms | candidate list, select_scratch_bytes(n)) lives in
// the library image instead of coming from cudaMallocAsync: one fixed array per GPU that loads
// this library, sized for CF_MEDIAN_NOALLOC_MAX_N elements (about 8.2 MB), held until the process
// exits.  See the declaration in curvefit_cuda.h for who needs this and what it costs.
namespace {

constexpr size_t kNoAllocScratchBytes =
    256 + (sizeof(unsigned int) << cf::kSelB0Single) + 2 * (sizeof(unsigned int) << cf::kSelWSingle)
    + sizeof(unsigned long long) * (size_t)CF_MEDIAN_NOALLOC_MAX_N;

// `g_median_scratch` is a device symbol.  Device code may name it; in HOST code the name is only a
// stand-in the runtime uses to look the symbol up -- it is not a device address and must never be
// handed to a kernel or to cudaMemsetAsync.  median_noalloc_scratch() asks the runtime for the real
// address on the CURRENT device (each device has its own copy of the array).
static __device__ __align__(256) unsigned char g_median_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoAllocMaxDevices];        // device address of g_median_scratch, per device ordinal

int median_noalloc_scratch(void** out)
{
    int dev = -1;
    CF_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kNoAllocMaxDevices) {
        cf::set_error("cf_median_device_noalloc: device ordinal beyond the per-device table (64)");
        return CF_DEVICE_ERROR;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {                  // first call on this device: resolve the symbol once
        void* p = nullptr;
        CF_CUDA(cudaGetSymbolAddress(&p, g_median_scratch));
        if (!p) { cf::set_error("cf_median_device_noalloc: scratch symbol has no device address"); return CF_DEVICE_ERROR; }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return CF_OK;
}

}  // namespace

extern "C" int32_t cf_median_device_noalloc(const void* d_x, int32_t n, void* d_z, void* stream)
{
    if (n < 1 || !d_x || !d_z) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (n > CF_MEDIAN_NOALLOC_MAX_N || select_scratch_bytes(n) > kNoAllocScratchBytes) {
        set_error("cf_median_device_noalloc: n = %d exceeds the built-in scratch (at most %d elements); "
                  "use cf_median_device", (int)n, (int)CF_MEDIAN_NOALLOC_MAX_N);
        return CF_ARRAY_SIZE_ERROR;
    }
    void* scratch = nullptr;
    CF_TRY(median_noalloc_scratch(&scratch));
    // select_median zeroes the state and both histograms itself (cudaMemsetAsync on `stream`), so
    // what the previous call left in the array does not matter; the candidate list is write-before-read.
    return select_median(static_cast<const double*>(d_x), n, static_cast<double*>(d_z), scratch,
                         (cudaStream_t)stream, nullptr);
}
