// fit.cu -- the local-fit kernel: one pass of lowest() for a block of output
// points (src/curvefit_regression.f90:513-586 as called from
// :671-673).  This kernel is > 99.9 % of smooth() and is FP64-pipe bound.
//
// Decomposition.  A CTA owns a tile of T = NT*P consecutive output points
// (P per thread, kept in registers) and streams the union of their windows
// through shared memory in double-buffered chunks of C elements.  Every window
// element is read by all threads as a shared-memory BROADCAST, so there is no
// reduction and each point's sums run over ascending j like the reference.
//
// Arithmetic.  The reference makes 5-6 dependent sweeps (weights, normalise,
// mean, variance, reweight, dot with y).  They collapse into one sweep over
// five moments of the tricube x robustness weights:
//     S0 = sum w      S1 = sum w u      S2 = sum w u^2
//     T0 = sum w y    T1 = sum w u y          u_j = (x_j - xc) / hc
// with (xc, hc) a per-CTA centre/scale, so the factors that depend only on j
//     rw, rw u, rw u^2, rw y, rw u y
// are formed ONCE per chunk while staging and every (point, element) pair costs
// 10 FP64-pipe instructions: q = fma(u, a_p, b_p) (= (x_j - xs)/h), q^2,
// t = 1 - q^2|q|, t^2, t^3, and five FMAs.  The three-way decisions of the
// reference (r <= h1 -> w = 1, r > h9 -> w = 0, sweep exit) are taken from the
// integer ranges produced by windows.cu, never from the rounded q, so the set
// of contributing elements is exactly the reference's.
#include "common.cuh"

namespace cf {

struct FitParams {
    const double* x;
    const double* y;
    const double* rw;   // nullptr on the first pass (userw = .false., :671)
    double* ys;
    Windows w;
    int32_t n, ns, begin, end;
};

__device__ __forceinline__ int warp_min(int v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_max(int v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_maxd(double v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

template <int NT, int P, int E>
__global__ void __launch_bounds__(NT)
fit_kernel(const FitParams prm)
{
    constexpr int C = NT * E;      // window elements per staged chunk
    constexpr int T = NT * P;      // output points per CTA
    constexpr int NW = NT / 32;

    __shared__ double2 sm[2][3][C];
    __shared__ int s_lo[NW], s_hi[NW];
    __shared__ double s_h[NW];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int32_t tile0 = prm.begin + (int32_t)blockIdx.x * T;
    const int32_t tile_end = min(tile0 + T, prm.end);
    const double* __restrict__ x = prm.x;
    const double* __restrict__ y = prm.y;
    const double* __restrict__ rw = prm.rw;

    // ---- per-point setup -----------------------------------------------------
    int32_t pi[P], lo9[P], hi9[P], lo1[P], hi1[P];
    double hp[P], xsp[P];
    bool valid[P];
    int wlo = INT32_MAX, whi = -1, flo = -1, fhi = INT32_MAX, clo = INT32_MAX, chi = -1;
    double hmax = 0.0;
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const int32_t i = tile0 + wid * (32 * P) + p * 32 + lane;
        pi[p] = i;
        // tie followers copy their leader's value later (robust.cu); a shard that
        // starts inside a tie group fits its first point in the leader's place
        bool v = i < tile_end;
        if (v) {
            const int32_t l = prm.w.leader[i];
            v = (l == i) || (i == prm.begin && l < i);
        }
        valid[p] = v;
        if (valid[p]) {
            const int32_t nl = prm.w.nleft[i];
            const double xs = x[i];
            const double hl = xs - x[nl], hr = x[nl + prm.ns - 1] - xs;   // :536
            hp[p] = hl > hr ? hl : hr;
            xsp[p] = xs;
            lo9[p] = prm.w.lo9[i]; hi9[p] = prm.w.hi9[i];
            lo1[p] = prm.w.lo1[i]; hi1[p] = prm.w.hi1[i];
            wlo = min(wlo, lo9[p]); whi = max(whi, hi9[p]);
            clo = min(clo, lo1[p]); chi = max(chi, hi1[p]);
            flo = max(flo, lo9[p]); fhi = min(fhi, hi9[p]);
            hmax = fmax(hmax, hp[p]);
        } else {
            hp[p] = 0.0; xsp[p] = 0.0;
            lo9[p] = 1; hi9[p] = 0; lo1[p] = 1; hi1[p] = 0;   // empty ranges
            flo = INT32_MAX; fhi = -1;                        // no fast chunks
        }
    }
    wlo = warp_min(wlo); whi = warp_max(whi);
    flo = warp_max(flo); fhi = warp_min(fhi);
    clo = warp_min(clo); chi = warp_max(chi);
    hmax = warp_maxd(hmax);
    if (lane == 0) { s_lo[wid] = wlo; s_hi[wid] = whi; s_h[wid] = hmax; }
    __syncthreads();
    int jlo = INT32_MAX, jhi = -1;
    double hc = 0.0;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        jlo = min(jlo, s_lo[w]); jhi = max(jhi, s_hi[w]); hc = fmax(hc, s_h[w]);
    }
    if (jlo > jhi) return;                     // tile holds tie followers only
    if (!(hc > 0.0)) hc = 1.0;
    const double sc = 1.0 / hc;
    const double xc = x[min(tile0 + T / 2, prm.n - 1)];

    double ap[P], bp[P];
    double S0[P], S1[P], S2[P], T0[P], T1[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const double invh = hp[p] > 0.0 ? 1.0 / hp[p] : 0.0;
        ap[p] = hc * invh;
        bp[p] = (xc - xsp[p]) * invh;
        S0[p] = S1[p] = S2[p] = T0[p] = T1[p] = 0.0;
    }

    // ---- staging -------------------------------------------------------------
    double rx[E], ry[E], rr[E];
    auto fetch = [&](int32_t c) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
            const int32_t j = c + e * NT + tid;
            const bool in = (j <= jhi);
            rx[e] = in ? x[j] : xc;
            ry[e] = in ? y[j] : 0.0;
            rr[e] = in ? (rw ? rw[j] : 1.0) : 0.0;
        }
    };
    auto stash = [&](int buf) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
            const int k = e * NT + tid;
            const double u = (rx[e] - xc) * sc;
            const double ru = rr[e] * u, ryy = rr[e] * ry[e];
            sm[buf][0][k] = make_double2(u, rr[e]);
            sm[buf][1][k] = make_double2(ru, ru * u);
            sm[buf][2][k] = make_double2(ryy, ryy * u);
        }
    };

    const int nchunks = (jhi - jlo) / C + 1;
    fetch(jlo);
    stash(0);
    __syncthreads();

    for (int kc = 0; kc < nchunks; ++kc) {
        const int32_t c = jlo + kc * C, cend = c + C - 1;
        const int buf = kc & 1;
        const bool more = kc + 1 < nchunks;
        if (more) fetch(c + C);

        const double2* __restrict__ p0 = sm[buf][0];
        const double2* __restrict__ p1 = sm[buf][1];
        const double2* __restrict__ p2 = sm[buf][2];

        if (cend < wlo || c > whi) {
            // nothing for this warp in the chunk
        } else if (c >= flo && cend <= fhi && (cend < clo || c > chi)) {
            // interior: every element lies in (h1, h9] of every point of the warp
#pragma unroll 2
            for (int k = 0; k < C; ++k) {
                const double2 v0 = p0[k], v1 = p1[k], v2 = p2[k];
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const double q = fma(v0.x, ap[p], bp[p]);
                    const double q2 = q * q;
                    const double t = fma(-q2, fabs(q), 1.0);
                    const double t2 = t * t;
                    const double w = t2 * t;
                    S0[p] = fma(w, v0.y, S0[p]);
                    S1[p] = fma(w, v1.x, S1[p]);
                    S2[p] = fma(w, v1.y, S2[p]);
                    T0[p] = fma(w, v2.x, T0[p]);
                    T1[p] = fma(w, v2.y, T1[p]);
                }
            }
        } else {
            // edge / core chunk: per-point integer range tests
#pragma unroll 2
            for (int k = 0; k < C; ++k) {
                const double2 v0 = p0[k], v1 = p1[k], v2 = p2[k];
                const int32_t j = c + k;
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    const double q = fma(v0.x, ap[p], bp[p]);
                    const double q2 = q * q;
                    const double t = fma(-q2, fabs(q), 1.0);
                    const double t2 = t * t;
                    double w = t2 * t;
                    w = (j >= lo1[p] && j <= hi1[p]) ? 1.0 : w;     // :547-551
                    w = (j >= lo9[p] && j <= hi9[p]) ? w : 0.0;     // :546, :553
                    S0[p] = fma(w, v0.y, S0[p]);
                    S1[p] = fma(w, v1.x, S1[p]);
                    S2[p] = fma(w, v1.y, S2[p]);
                    T0[p] = fma(w, v2.x, T0[p]);
                    T1[p] = fma(w, v2.y, T1[p]);
                }
            }
        }
        if (more) stash(buf ^ 1);
        __syncthreads();
    }

    // ---- epilogue: local linear fit evaluated at xs (:559-584) ------------------
    const double range = x[prm.n - 1] - x[0];                              // :535
#pragma unroll
    for (int p = 0; p < P; ++p) {
        if (!valid[p]) continue;
        const int32_t i = pi[p];
        double out;
        if (!(S0[p] > 0.0)) {
            out = y[i];                                   // ok = .false., :559, :673
        } else {
            const double inv = 1.0 / S0[p];
            const double ym = T0[p] * inv;                // weighted mean of y
            out = ym;
            if (hp[p] > 0.0) {                            // :564
                const double um = S1[p] * inv;
                double cu = fma(-um, um, S2[p] * inv);    // weighted variance of u
                cu = cu > 0.0 ? cu : 0.0;
                if (sqrt(cu) * hc > 0.001 * range) {      // :573 (c = cu * hc^2)
                    const double us = (xsp[p] - xc) * sc;
                    const double cov = fma(-um, ym, T1[p] * inv);
                    out = fma(cov / cu, us - um, ym);
                }
            }
        }
        prm.ys[i] = out;
    }
}

int launch_fit(cf_lowess* h, bool use_rw)
{
    if (h->n < 2 || h->end <= h->begin) return CF_OK;
    constexpr int NT = 128, P = 4, E = 2;
    FitParams prm;
    prm.x = h->x; prm.y = h->y; prm.rw = use_rw ? h->rw : nullptr; prm.ys = h->ys;
    prm.w = h->win; prm.n = h->n; prm.ns = h->ns; prm.begin = h->begin; prm.end = h->end;
    const int T = NT * P;
    const int tiles = (h->end - h->begin + T - 1) / T;
    fit_kernel<NT, P, E><<<tiles, NT, 0, h->stream>>>(prm);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    h->fit_launches++;
    return CF_OK;
}

}  // namespace cf
