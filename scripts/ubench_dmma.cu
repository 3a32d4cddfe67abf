// Development micro-benchmark: FP64 tensor (DMMA) rate on B200 and whether it overlaps
// with the FP64 vector pipe.
#include <cstdio>
#include <cuda_runtime.h>
constexpr int ITERS = 1 << 14;

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma1688(double& c0, double& c1, double& c2, double& c3,
                                         double a0, double a1, double a2, double a3, double b0, double b1)
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                 : "+d"(c0), "+d"(c1), "+d"(c2), "+d"(c3) : "d"(a0), "d"(a1), "d"(a2), "d"(a3), "d"(b0), "d"(b1));
}

template <int NACC>
__global__ void __launch_bounds__(128) k_dmma(double* out, const double* in)
{
    double c[NACC][2];
    for (int k = 0; k < NACC; ++k) c[k][0] = c[k][1] = 0.0;
    double a = in[threadIdx.x], b = in[threadIdx.x + 128];
    for (int it = 0; it < ITERS; ++it)
#pragma unroll
        for (int k = 0; k < NACC; ++k) dmma884(c[k][0], c[k][1], a, b);
    double s = 0; for (int k = 0; k < NACC; ++k) s += c[k][0] + c[k][1];
    if (s == 123.456) out[0] = s;
}

template <int NACC>
__global__ void __launch_bounds__(128) k_dmma16(double* out, const double* in)
{
    double c[NACC][4];
    for (int k = 0; k < NACC; ++k) c[k][0] = c[k][1] = c[k][2] = c[k][3] = 0.0;
    double a0 = in[threadIdx.x], a1 = in[threadIdx.x + 1], a2 = in[threadIdx.x + 2], a3 = in[threadIdx.x + 3];
    double b0 = in[threadIdx.x + 128], b1 = in[threadIdx.x + 129];
    for (int it = 0; it < ITERS; ++it)
#pragma unroll
        for (int k = 0; k < NACC; ++k) dmma1688(c[k][0], c[k][1], c[k][2], c[k][3], a0, a1, a2, a3, b0, b1);
    double s = 0; for (int k = 0; k < NACC; ++k) s += c[k][0] + c[k][1] + c[k][2] + c[k][3];
    if (s == 123.456) out[0] = s;
}

// NF DFMAs (2 vector operands + uniform) per DMMA
template <int NACC, int NF>
__global__ void __launch_bounds__(128) k_mix(double* out, const double* in, double ua, double ub)
{
    double c[NACC][2], v[8];
    for (int k = 0; k < NACC; ++k) c[k][0] = c[k][1] = 0.0;
    for (int k = 0; k < 8; ++k) v[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    double a = in[threadIdx.x], b = in[threadIdx.x + 128];
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < NACC; ++k) {
            dmma884(c[k][0], c[k][1], a, b);
#pragma unroll
            for (int f = 0; f < NF; ++f) v[(k * NF + f) & 7] = fma(v[(k * NF + f) & 7], ua, ub);
        }
    }
    double s = 0; for (int k = 0; k < NACC; ++k) s += c[k][0] + c[k][1];
    for (int k = 0; k < 8; ++k) s += v[k];
    if (s == 123.456) out[0] = s;
}

template <typename F> float timeit(F f)
{
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main()
{
    double *out, *in; cudaMalloc(&out, 8); cudaMalloc(&in, 8 * 4096);
    double h[4096]; for (int i = 0; i < 4096; ++i) h[i] = 0.001 * (i % 97) + 0.5;
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    for (int wps : {1, 2, 4}) {
        const int blocks = sms * wps;
        const double warps = (double)blocks * 4;
        auto tf = [&](double fma_per_warp, float ms) { return 2.0 * fma_per_warp * warps / (ms * 1e-3) / 1e12; };
        float m1 = timeit([&] { k_dmma<8><<<blocks, 128>>>(out, in); });
        float m2 = timeit([&] { k_dmma16<4><<<blocks, 128>>>(out, in); });
        float m3 = timeit([&] { k_mix<8, 4><<<blocks, 128>>>(out, in, 0.999999, 1e-6); });
        float m4 = timeit([&] { k_mix<8, 8><<<blocks, 128>>>(out, in, 0.999999, 1e-6); });
        printf("warps/SMSP=%d: dmma884 %.1f TF | dmma16x8x8 %.1f TF | mix(1 dmma + 4 dfma): dmma %.1f + dfma %.1f TF | mix(1+8): dmma %.1f + dfma %.1f TF\n",
               wps, tf(256.0 * 8 * ITERS, m1), tf(1024.0 * 4 * ITERS, m2),
               tf(256.0 * 8 * ITERS, m3), tf(32.0 * 4 * 8 * ITERS, m3),
               tf(256.0 * 8 * ITERS, m4), tf(32.0 * 8 * 8 * ITERS, m4));
    }
    return 0;
}
