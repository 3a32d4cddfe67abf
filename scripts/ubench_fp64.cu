// Development micro-benchmark: what FP64-pipe rate can the fit kernel's instruction
// mix reach when nothing but registers is involved?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench scripts/ubench_fp64.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 1 << 14;

__global__ void __launch_bounds__(128) v0_uniform(double* out, double a, double b)
{
    double v[8];
    for (int k = 0; k < 8; ++k) v[k] = 1.0 + 1e-9 * (threadIdx.x + k);
    for (int it = 0; it < ITERS; ++it)
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = fma(v[k], a, b);
    double s = 0; for (int k = 0; k < 8; ++k) s += v[k];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(128) v1_threereg(double* out, const double* in)
{
    double v[8], a[8], b[8];
    for (int k = 0; k < 8; ++k) { v[k] = 1.0 + 1e-9 * (threadIdx.x + k); a[k] = in[k + threadIdx.x]; b[k] = in[8 + k + threadIdx.x]; }
    for (int it = 0; it < ITERS; ++it)
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = fma(v[k], a[k], b[k]);
    double s = 0; for (int k = 0; k < 8; ++k) s += v[k];
    if (s == 123.456) out[0] = s;
}

// the fit kernel's fast-path body for P points, element values held in registers and
// perturbed by an integer op per iteration so that nothing can be hoisted
template <int P>
__global__ void __launch_bounds__(128) v2_body(double* out, const double* in)
{
    double ap[P], bp[P], S0[P], S1[P], S2[P], T0[P], T1[P];
    for (int p = 0; p < P; ++p) { ap[p] = in[p + threadIdx.x]; bp[p] = in[8 + p]; S0[p] = S1[p] = S2[p] = T0[p] = T1[p] = 0; }
    double e[6];
    for (int k = 0; k < 6; ++k) e[k] = in[16 + k];
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            long long bits = __double_as_longlong(e[k]);
            bits ^= (long long)(it & 3);
            e[k] = __longlong_as_double(bits);
        }
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const double q = fma(e[0], ap[p], bp[p]);
            const double q2 = q * q;
            const double t = fma(-q2, fabs(q), 1.0);
            const double t2 = t * t;
            const double w = t2 * t;
            S0[p] = fma(w, e[1], S0[p]);
            S1[p] = fma(w, e[2], S1[p]);
            S2[p] = fma(w, e[3], S2[p]);
            T0[p] = fma(w, e[4], T0[p]);
            T1[p] = fma(w, e[5], T1[p]);
        }
    }
    double s = 0; for (int p = 0; p < P; ++p) s += S0[p] + S1[p] + S2[p] + T0[p] + T1[p];
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
    for (int warps_per_smsp : {1, 2, 3, 4}) {
        const int blocks = sms * warps_per_smsp;   // 128-thread CTAs: 4 warps = 1 per SMSP
        auto rate = [&](double instr_per_thread, float ms) { return instr_per_thread * blocks * 128 / (ms * 1e-3) / (64.0 * sms) / 1e6; };
        float m0 = timeit([&] { v0_uniform<<<blocks, 128>>>(out, 0.999999, 1e-6); });
        float m1 = timeit([&] { v1_threereg<<<blocks, 128>>>(out, in); });
        float m2 = timeit([&] { v2_body<2><<<blocks, 128>>>(out, in); });
        float m4 = timeit([&] { v2_body<4><<<blocks, 128>>>(out, in); });
        float m8 = timeit([&] { v2_body<8><<<blocks, 128>>>(out, in); });
        printf("warps/SMSP=%d  equiv MHz (1965 = full pipe): uniform %.0f  threereg %.0f  body P=2 %.0f  P=4 %.0f  P=8 %.0f\n",
               warps_per_smsp, rate(8.0 * ITERS, m0), rate(8.0 * ITERS, m1), rate(20.0 * ITERS, m2),
               rate(40.0 * ITERS, m4), rate(80.0 * ITERS, m8));
    }
    return 0;
}
