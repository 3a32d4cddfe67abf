// peak.cu -- FP64 FMA-pipe peak microbenchmark.  MEASURED_PEAKS.json carries no
// FP64 figure, and the LOWESS fit kernel is FP64-pipe bound, so the roofline
// denominator is measured here: 8 independent DFMA chains per thread, 8 warps
// per SM sub-partition, no memory traffic.
#include "common.cuh"

namespace cf {

constexpr int kPeakChains = 8;
constexpr int kPeakInner = 4096;

__global__ void __launch_bounds__(256)
dfma_peak_kernel(double* out, double a, double b, int reps)
{
    double v[kPeakChains];
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) v[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    for (int r = 0; r < reps; ++r) {
#pragma unroll 16
        for (int it = 0; it < kPeakInner; ++it) {
#pragma unroll
            for (int k = 0; k < kPeakChains; ++k) v[k] = fma(v[k], a, b);
        }
    }
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) s += v[k];
    if (s == 123.456) out[0] = s;   // keeps the chains alive
}

int fp64_peak(int device, double ms_target, double* tflops, double* mhz)
{
    if (device >= 0) CF_CUDA(cudaSetDevice(device));
    int dev = 0, sms = 0;
    CF_CUDA(cudaGetDevice(&dev));
    CF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    double* d_out = nullptr;
    CF_CUDA(cudaMalloc(&d_out, sizeof(double)));
    cudaEvent_t e0, e1;
    CF_CUDA(cudaEventCreate(&e0));
    CF_CUDA(cudaEventCreate(&e1));
    const int blocks = sms * 8, threads = 256;
    int reps = 4;
    // warm-up, then scale reps so that one launch lasts about ms_target
    dfma_peak_kernel<<<blocks, threads>>>(d_out, 0.999999, 1e-6, reps);
    CF_CUDA(cudaDeviceSynchronize());
    float ms = 0.f;
    for (int attempt = 0; attempt < 3; ++attempt) {
        CF_CUDA(cudaEventRecord(e0));
        dfma_peak_kernel<<<blocks, threads>>>(d_out, 0.999999, 1e-6, reps);
        CF_CUDA(cudaEventRecord(e1));
        CF_CUDA(cudaEventSynchronize(e1));
        CF_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (ms >= 0.5 * ms_target || reps > (1 << 24)) break;
        const double scale = ms > 0.01 ? ms_target / ms : 16.0;
        reps = (int)(reps * scale) + 1;
    }
    const double fmas = (double)blocks * threads * kPeakChains * (double)kPeakInner * reps;
    if (tflops) *tflops = 2.0 * fmas / (ms * 1e-3) / 1e12;
    // 64 DFMA lanes per SM per clock
    if (mhz) *mhz = fmas / (ms * 1e-3) / (64.0 * sms) / 1e6;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    return CF_OK;
}

}  // namespace cf
