// fit.cu -- the local-fit kernel: one pass of lowest() for a block of output
// points (src/curvefit_regression.f90:513-586 as called from
// :671-673).  This kernel is > 99.9 % of smooth() and is FP64-pipe bound.
//
// Decomposition.  A tile is T = NT*P consecutive output points (P per thread,
// accumulators in registers).  A CTA streams (a K-th part of) the union of the
// tile's windows through shared memory in double-buffered chunks of C elements;
// every window element is read by all threads as a shared-memory BROADCAST, so
// there is no reduction and each point's sums run over ascending j like the
// reference.  When the tile count does not fill the GPU evenly, every tile's
// window is cut into K parts handled by K CTAs (partial moments to scratch,
// combined in part order by fit_finalize_kernel), so the grid is many waves
// deep and the tail is short.
//
// Arithmetic.  The reference makes 5-6 dependent sweeps (weights, normalise,
// mean, variance, reweight, dot with y).  They collapse into one sweep over
// five moments of the tricube x robustness weights:
//     S0 = sum w      S1 = sum w u      S2 = sum w u^2
//     T0 = sum w y    T1 = sum w u y
// FAST path: u_j = (x_j - xc)/hc with a per-tile centre/scale, so the factors
// that depend only on j (rw, rw u, rw u^2, rw y, rw u y) are formed ONCE per
// chunk while staging and every (point, element) pair costs 10 FP64-pipe
// instructions: q = fma(u, a_p, b_p) (= (x_j - xs)/h), q^2, t = 1 - q^2|q|,
// t^2, t^3, five FMAs.  EXACT path (tiles whose bandwidths vary so much that a
// shared centre/scale would lose digits: window width / min h > 16): u = q is
// formed per point from d = x_j - xs exactly as the reference forms r, 13
// instructions per pair.  The three-way decisions of the reference (r <= h1 ->
// w = 1, r > h9 -> w = 0, sweep exit) are taken from the integer ranges of
// windows.cu, never from the rounded q, so the set of contributing elements is
// exactly the reference's.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace cf {

struct FitParams {
    const double* x;
    const double* y;
    const double* rw;   // nullptr on the first pass (userw = .false., :671)
    double* ys;
    double* part;       // [tile][K][5][T] partial moments when K > 1
    Windows w;
    int32_t n, ns, begin, end, ksplit;
    int32_t tiles;      // tiles per series (launched)
    int32_t tile_first, tile_stride;   // launched tile t is tile tile_first + t * tile_stride of the shard
    const int32_t* fidx;  // delta > 0: output slot -> point (delta.cu); then begin = 0, end = #fitted
};

// output slot -> index of the point fitted there
__device__ __forceinline__ int32_t slot_point(const FitParams& prm, int32_t slot)
{
    return prm.fidx ? __ldg(prm.fidx + slot) : slot;
}

// batched mode: series blockIdx.y occupies [series * n, (series + 1) * n) of every array
__device__ __forceinline__ FitParams for_series(const FitParams& in, int series)
{
    FitParams p = in;
    const size_t off = (size_t)series * (size_t)in.n;
    p.x += off; p.y += off; p.ys += off;
    if (p.rw) p.rw += off;
    p.w.nleft += off; p.w.lo9 += off; p.w.hi9 += off;
    p.w.lo1 += off; p.w.hi1 += off; p.w.leader += off;
    if (p.part) p.part += (size_t)series * (size_t)in.tiles * in.ksplit * 5 * 512;
    return p;
}

constexpr double kExactRatio = 16.0;

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
__device__ __forceinline__ double warp_mind(double v)
{
#pragma unroll
    for (int o = 16; o; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Tile-level facts shared by the fit and finalize kernels.
template <int NT, int P>
struct Tile {
    static constexpr int T = NT * P;
    static constexpr int NW = NT / 32;
    // per thread
    bool valid[P];
    double hp[P], xsp[P];
    // per warp
    int wlo, whi, flo, fhi, clo, chi;
    // per CTA
    int jlo, jhi;
    double xc, hc, sc;
    bool exact;
};

struct TileShared {
    int lo[8], hi[8];
    double hmax[8], hmin[8];
};

// Reads the windows of the tile's points, publishes the per-point integer ranges
// to shared memory (rng[4][T]) and reduces the tile-level quantities.
template <int NT, int P>
__device__ __forceinline__ void tile_setup(const FitParams& prm, int32_t tile0, Tile<NT, P>& t,
                                           int32_t (*rng)[NT * P], TileShared& ts)
{
    constexpr int T = NT * P, NW = NT / 32;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int32_t tile_end = min(tile0 + T, prm.end);
    const double* __restrict__ x = prm.x;
    int wlo = INT32_MAX, whi = -1, flo = -1, fhi = INT32_MAX, clo = INT32_MAX, chi = -1;
    double hmax = 0.0, hmin = 1.7976931348623157e308;
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const int loc = wid * (32 * P) + p * 32 + lane;
        // tie followers copy their leader's value later (robust.cu); a shard that
        // starts inside a tie group fits its first point in the leader's place.
        // With a fitted-point list every slot is fitted.
        bool v = tile0 + loc < tile_end;
        const int32_t i = v ? slot_point(prm, tile0 + loc) : 0;
        if (v && !prm.fidx) {
            const int32_t l = prm.w.leader[i];
            v = (l == i) || (i == prm.begin && l < i);
        }
        t.valid[p] = v;
        int l9 = 1, r9 = 0, l1 = 1, r1 = 0;       // empty ranges for idle slots
        if (v) {
            const int32_t nl = prm.w.nleft[i];
            const double xs = x[i];
            const double hl = xs - x[nl], hr = x[nl + prm.ns - 1] - xs;   // :536
            const double h = hl > hr ? hl : hr;
            t.hp[p] = h;
            t.xsp[p] = xs;
            l9 = prm.w.lo9[i]; r9 = prm.w.hi9[i];
            l1 = prm.w.lo1[i]; r1 = prm.w.hi1[i];
            wlo = min(wlo, l9); whi = max(whi, r9);
            clo = min(clo, l1); chi = max(chi, r1);
            flo = max(flo, l9); fhi = min(fhi, r9);
            hmax = fmax(hmax, h);
            if (h > 0.0) hmin = fmin(hmin, h);
        } else {
            // an idle slot (tie follower, past the tile's end) constrains nothing: its sums are
            // never stored, so the warp's predicate-free chunks may run over it with any weights
            // (it used to switch them off for the whole warp -- every warp of a series with ties)
            t.hp[p] = 0.0; t.xsp[p] = 0.0;
        }
        rng[0][loc] = l9; rng[1][loc] = r9; rng[2][loc] = l1; rng[3][loc] = r1;
    }
    t.wlo = warp_min(wlo); t.whi = warp_max(whi);
    t.flo = warp_max(flo); t.fhi = warp_min(fhi);
    t.clo = warp_min(clo); t.chi = warp_max(chi);
    hmax = warp_maxd(hmax); hmin = warp_mind(hmin);
    if (lane == 0) { ts.lo[wid] = t.wlo; ts.hi[wid] = t.whi; ts.hmax[wid] = hmax; ts.hmin[wid] = hmin; }
    __syncthreads();
    int jlo = INT32_MAX, jhi = -1;
    double hc = 0.0, hm = 1.7976931348623157e308;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        jlo = min(jlo, ts.lo[w]); jhi = max(jhi, ts.hi[w]);
        hc = fmax(hc, ts.hmax[w]); hm = fmin(hm, ts.hmin[w]);
    }
    t.jlo = jlo; t.jhi = jhi;
    t.exact = false;
    if (jlo <= jhi) {
        // a shared centre/scale loses digits when some bandwidth is tiny next to
        // the width of the tile's window union
        const double width = x[jhi] - x[jlo];
        t.exact = hc > 0.0 && !(width <= kExactRatio * hm);
    }
    if (!(hc > 0.0)) hc = 1.0;
    t.hc = hc;
    t.sc = 1.0 / hc;
    t.xc = x[slot_point(prm, min(tile0 + T / 2, (prm.fidx ? prm.end : prm.n) - 1))];
}

template <int NT, int P, int E>
constexpr size_t fit_smem_bytes()
{
    return sizeof(double2) * 2 * 3 * (size_t)(NT * E) + sizeof(int32_t) * 4 * (size_t)(NT * P);
}

// CL (opt-in, see launch_variant): the K <= 8 CTAs of a tile form a thread-block CLUSTER and
// add their partial moments through distributed shared memory in part order.  Default: the
// parts go through prm.part and fit_finalize_kernel.
template <int NT, int P, int E, int MINB, int U, bool CL>
__global__ void __launch_bounds__(NT, MINB)
fit_kernel(const FitParams prm_in)
{
    const FitParams prm = for_series(prm_in, blockIdx.y);
    constexpr int C = NT * E;      // window elements per staged chunk
    constexpr int T = NT * P;      // output points per tile

    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2 (*sm)[3][C] = reinterpret_cast<double2 (*)[3][C]>(smem_raw);
    int32_t (*rng)[T] = reinterpret_cast<int32_t (*)[T]>(smem_raw + sizeof(double2) * 2 * 3 * C);
    __shared__ TileShared ts;

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int K = prm.ksplit;
    const int tile = blockIdx.x / K, part = blockIdx.x - tile * K;
    const int32_t tile0 = prm.begin + (prm.tile_first + (int32_t)tile * prm.tile_stride) * T;
    const double* __restrict__ x = prm.x;
    const double* __restrict__ y = prm.y;
    const double* __restrict__ rw = prm.rw;

    Tile<NT, P> t;
    tile_setup<NT, P>(prm, tile0, t, rng, ts);
    const int jlo = t.jlo, jhi = t.jhi;
    if (!CL && jlo > jhi) return;              // tile holds tie followers only (a cluster stays for its barriers)
    const double xc = t.xc, sc = t.sc;
    const bool exact = t.exact;

    // per-point constants: fast q = fma(u, ap, bp); exact q = (x_j - bp) * ap
    double ap[P], bp[P];
    double S0[P], S1[P], S2[P], T0[P], T1[P];
#pragma unroll
    for (int p = 0; p < P; ++p) {
        const double invh = t.hp[p] > 0.0 ? 1.0 / t.hp[p] : 0.0;
        ap[p] = exact ? invh : t.hc * invh;
        bp[p] = exact ? t.xsp[p] : (xc - t.xsp[p]) * invh;
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
            if (exact) {
                sm[buf][0][k] = make_double2(rx[e], rr[e]);
                sm[buf][1][k] = make_double2(ry[e], 0.0);
            } else {
                const double u = (rx[e] - xc) * sc;
                const double ru = rr[e] * u, ryy = rr[e] * ry[e];
                sm[buf][0][k] = make_double2(u, rr[e]);
                sm[buf][1][k] = make_double2(ru, ru * u);
                sm[buf][2][k] = make_double2(ryy, ryy * u);
            }
        }
    };

    const int nch = jlo > jhi ? 0 : (jhi - jlo) / C + 1;
    const int k0 = (int)(((long long)nch * part) / K);
    const int k1 = (int)(((long long)nch * (part + 1)) / K);
    if (k0 < k1) {
        fetch(jlo + k0 * C);
        stash(k0 & 1);
    }
    __syncthreads();

    for (int kc = k0; kc < k1; ++kc) {
        const int32_t c = jlo + kc * C, cend = c + C - 1;
        const int buf = kc & 1;
        const bool more = kc + 1 < k1;
        if (more) fetch(c + C);

        const double2* __restrict__ p0 = sm[buf][0];
        const double2* __restrict__ p1 = sm[buf][1];
        const double2* __restrict__ p2 = sm[buf][2];

        if (cend < t.wlo || c > t.whi) {
            // nothing for this warp in the chunk
        } else if (exact) {
            // per-point centring, 13 FP64 instructions per pair, integer range tests
#pragma unroll
            for (int p = 0; p < P; ++p) {
                const int loc = wid * (32 * P) + p * 32 + lane;
                const int l9 = rng[0][loc], r9 = rng[1][loc], l1 = rng[2][loc], r1 = rng[3][loc];
                double s0 = S0[p], s1 = S1[p], s2 = S2[p], u0 = T0[p], u1 = T1[p];
#pragma unroll 4
                for (int k = 0; k < C; ++k) {
                    const double2 v0 = p0[k], v1 = p1[k];
                    const int32_t j = c + k;
                    const double q = (v0.x - bp[p]) * ap[p];
                    const double q2 = q * q;
                    const double tt = fma(-q2, fabs(q), 1.0);
                    double w = (tt * tt) * tt;
                    w = (j >= l1 && j <= r1) ? 1.0 : w;     // :547-551
                    w = (j >= l9 && j <= r9) ? w : 0.0;     // :546, :553
                    const double wr = w * v0.y;
                    const double wq = wr * q;
                    s0 += wr;
                    s1 += wq;
                    s2 = fma(wq, q, s2);
                    u0 = fma(wr, v1.x, u0);
                    u1 = fma(wq, v1.x, u1);
                }
                S0[p] = s0; S1[p] = s1; S2[p] = s2; T0[p] = u0; T1[p] = u1;
            }
        } else {
            // Sub-blocks of SB elements are classified per warp: "interior" ones lie
            // strictly inside (h1, h9] of all 128 points of the warp and need no
            // predicate; edge / core ones test the integer ranges per point.
            constexpr int SB = 32;
            for (int sb = 0; sb < C; sb += SB) {
                const int32_t c0 = c + sb, c1 = c0 + SB - 1;
                if (c1 < t.wlo || c0 > t.whi) continue;
                if (c0 >= t.flo && c1 <= t.fhi && (c1 < t.clo || c0 > t.chi)) {
                    // software-pipelined broadcast loads: element k+1 is fetched while
                    // element k's 40 FP64 instructions issue
                    double2 n0 = p0[sb], n1 = p1[sb], n2 = p2[sb];
#pragma unroll U
                    for (int k = sb; k < sb + SB; ++k) {
                        const double2 v0 = n0, v1 = n1, v2 = n2;
                        n0 = p0[k + 1]; n1 = p1[k + 1]; n2 = p2[k + 1];
#pragma unroll
                        for (int p = 0; p < P; ++p) {
                            const double q = fma(v0.x, ap[p], bp[p]);
                            const double q2 = q * q;
                            const double tt = fma(-q2, fabs(q), 1.0);
                            const double t2 = tt * tt;
                            const double w = t2 * tt;
                            S0[p] = fma(w, v0.y, S0[p]);
                            S1[p] = fma(w, v1.x, S1[p]);
                            S2[p] = fma(w, v1.y, S2[p]);
                            T0[p] = fma(w, v2.x, T0[p]);
                            T1[p] = fma(w, v2.y, T1[p]);
                        }
                    }
                } else {
                    int l9[P], r9[P], l1[P], r1[P];
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        const int loc = wid * (32 * P) + p * 32 + lane;
                        l9[p] = rng[0][loc]; r9[p] = rng[1][loc];
                        l1[p] = rng[2][loc]; r1[p] = rng[3][loc];
                    }
#pragma unroll 2
                    for (int k = sb; k < sb + SB; ++k) {
                        const double2 v0 = p0[k], v1 = p1[k], v2 = p2[k];
                        const int32_t j = c + k;
#pragma unroll
                        for (int p = 0; p < P; ++p) {
                            const double q = fma(v0.x, ap[p], bp[p]);
                            const double q2 = q * q;
                            const double tt = fma(-q2, fabs(q), 1.0);
                            double w = (tt * tt) * tt;
                            w = (j >= l1[p] && j <= r1[p]) ? 1.0 : w;     // :547-551
                            w = (j >= l9[p] && j <= r9[p]) ? w : 0.0;     // :546, :553
                            S0[p] = fma(w, v0.y, S0[p]);
                            S1[p] = fma(w, v1.x, S1[p]);
                            S2[p] = fma(w, v1.y, S2[p]);
                            T0[p] = fma(w, v2.x, T0[p]);
                            T1[p] = fma(w, v2.y, T1[p]);
                        }
                    }
                }
            }
        }
        if (more) stash(buf ^ 1);
        __syncthreads();
    }

    if (CL) {
        // Every CTA parks its partial moments in its own shared memory (the staging
        // buffers are free now); CTA 0 of the cluster adds the K parts in part order
        // (= ascending j, the same order as fit_finalize_kernel) over DSMEM.
        cg::cluster_group cluster = cg::this_cluster();
        double* mine = reinterpret_cast<double*>(smem_raw);          // [5][T] <= 2*3*C double2
        static_assert(sizeof(double) * 5 * T <= sizeof(double2) * 2 * 3 * C, "partials must fit the staging buffers");
        __syncthreads();
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int loc = wid * (32 * P) + p * 32 + lane;
            mine[0 * T + loc] = S0[p];
            mine[1 * T + loc] = S1[p];
            mine[2 * T + loc] = S2[p];
            mine[3 * T + loc] = T0[p];
            mine[4 * T + loc] = T1[p];
        }
        cluster.sync();
        if (part == 0 && jlo <= jhi) {
            const double range = x[prm.n - 1] - x[0];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                if (!t.valid[p]) continue;
                const int loc = wid * (32 * P) + p * 32 + lane;
                double m[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
                for (int k = 0; k < K; ++k) {
                    const double* src = cluster.map_shared_rank(mine, k);
#pragma unroll
                    for (int q = 0; q < 5; ++q) m[q] += src[q * T + loc];
                }
                const int32_t i = slot_point(prm, tile0 + loc);
                const double scale = exact ? t.hp[p] : t.hc;
                const double us = exact ? 0.0 : (t.xsp[p] - xc) * sc;
                prm.ys[i] = fit_value(m[0], m[1], m[2], m[3], m[4], y[i], t.hp[p], scale, us, range);
            }
        }
        cluster.sync();                       // nobody leaves while CTA 0 still reads its memory
        return;
    }
    if (K > 1) {
        // partial moments; fit_finalize_kernel adds the K parts in part order
        double* __restrict__ dst = prm.part + ((size_t)tile * K + part) * (5 * (size_t)T);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int loc = wid * (32 * P) + p * 32 + lane;
            dst[0 * T + loc] = S0[p];
            dst[1 * T + loc] = S1[p];
            dst[2 * T + loc] = S2[p];
            dst[3 * T + loc] = T0[p];
            dst[4 * T + loc] = T1[p];
        }
        return;
    }

    const double range = x[prm.n - 1] - x[0];                              // :535
#pragma unroll
    for (int p = 0; p < P; ++p) {
        if (!t.valid[p]) continue;
        const int32_t i = slot_point(prm, tile0 + wid * (32 * P) + p * 32 + lane);
        const double scale = exact ? t.hp[p] : t.hc;
        const double us = exact ? 0.0 : (t.xsp[p] - xc) * sc;
        prm.ys[i] = fit_value(S0[p], S1[p], S2[p], T0[p], T1[p], y[i], t.hp[p], scale, us, range);
    }
}

// Combines the K partial moment sets of every tile (ascending part = ascending j)
// and evaluates the fit.  Launched with the fit kernel's CTA shape so that the
// tile-level quantities (xc, hc, exact) are recomputed identically.
template <int NT, int P>
__global__ void __launch_bounds__(NT)
fit_finalize_kernel(const FitParams prm_in)
{
    const FitParams prm = for_series(prm_in, blockIdx.y);
    constexpr int T = NT * P;
    static_assert(T == 512, "for_series() assumes 512-point tiles");
    __shared__ int32_t rng[4][T];
    __shared__ TileShared ts;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int K = prm.ksplit;
    const int tile = blockIdx.x;
    const int32_t tile0 = prm.begin + (prm.tile_first + (int32_t)tile * prm.tile_stride) * T;
    Tile<NT, P> t;
    tile_setup<NT, P>(prm, tile0, t, rng, ts);
    if (t.jlo > t.jhi) return;
    const double range = prm.x[prm.n - 1] - prm.x[0];
    const double* __restrict__ src = prm.part + (size_t)tile * K * (5 * (size_t)T);
#pragma unroll
    for (int p = 0; p < P; ++p) {
        if (!t.valid[p]) continue;
        const int loc = wid * (32 * P) + p * 32 + lane;
        double m[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
        for (int k = 0; k < K; ++k) {
#pragma unroll
            for (int q = 0; q < 5; ++q) m[q] += src[((size_t)k * 5 + q) * T + loc];
        }
        const int32_t i = slot_point(prm, tile0 + loc);
        const double scale = t.exact ? t.hp[p] : t.hc;
        const double us = t.exact ? 0.0 : (t.xsp[p] - t.xc) * t.sc;
        prm.ys[i] = fit_value(m[0], m[1], m[2], m[3], m[4], prm.y[i], t.hp[p], scale, us, range);
    }
}

// ---- launcher ---------------------------------------------------------------------------
namespace {

template <int NT, int P, int E, int MINB, int U>
int launch_variant(cf_lowess* h, FitParams& prm, int sms)
{
    constexpr int T = NT * P, C = NT * E;
    constexpr size_t smem = fit_smem_bytes<NT, P, E>();
    auto kern = fit_kernel<NT, P, E, MINB, U, false>;
    auto kern_cl = fit_kernel<NT, P, E, MINB, U, true>;
    CF_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CF_CUDA(cudaFuncSetAttribute(kern_cl, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int tiles = (prm.end - prm.begin + T - 1) / T;
    const bool sampled = h->tile_count > 0;      // cf_lowess_fit_sample: a strided subset of the tiles
    prm.tile_first = 0;
    prm.tile_stride = 1;
    if (sampled) {
        prm.tile_first = h->tile_first;
        prm.tile_stride = h->tile_stride;
        const int avail = tiles > h->tile_first ? (tiles - h->tile_first + h->tile_stride - 1) / h->tile_stride : 0;
        tiles = h->tile_count < avail ? h->tile_count : avail;
        if (tiles <= 0) return CF_OK;
    }
    const long long all_tiles = (long long)tiles * h->nseries;
    prm.tiles = tiles;
    // K-way split of every tile's window when the tiles alone give too few waves
    int occ = MINB;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, NT, smem);
    if (occ < 1) occ = 1;
    const long long slots = (long long)sms * occ;
    int K = 1;
    if (h->ksplit_override > 0) {
        K = h->ksplit_override;
    } else if (all_tiles < 48 * slots) {
        K = (int)((32 * slots + all_tiles - 1) / all_tiles);
    }
    const int maxk = (h->ns + T) / (4 * C) > 1 ? (h->ns + T) / (4 * C) : 1;   // >= 4 chunks per part
    if (K > maxk) K = maxk;
    if (K > 64) K = 64;
    if (K < 1) K = 1;
    if (sampled) K = h->tile_ksplit > 0 ? h->tile_ksplit : 1;   // the split the FULL launch would use
    // CF_FIT_VARIANT=4 (development): K <= 8 parts as one thread-block cluster that adds its
    // partial moments through DSMEM -- no scratch in HBM, no second kernel.  Measured on B200
    // at n = 1e6, ns = 5e4: 34.7 ms per pass against 32.7 ms for the scratch path (cluster
    // placement costs more than the 0.4 GB of partial-moment traffic it saves), so it is off.
    if (h->fit_variant == 4 && K > 8) K = 8;
    const bool use_cluster = h->fit_variant == 4 && K > 1 && K <= 8;
    prm.ksplit = K;
    prm.part = nullptr;
    if (use_cluster) {
        if (h->reserve_only) return CF_OK;                     // DSMEM variant: no scratch
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(tiles * K, h->nseries);
        cfg.blockDim = dim3(NT);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = h->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = K;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        CF_CUDA(cudaLaunchKernelEx(&cfg, kern_cl, prm));
        h->launches++;
        h->fit_launches++;
        return CF_OK;
    }
    if (K > 1) {
        const size_t need = (size_t)all_tiles * K * 5 * T * sizeof(double);
        if (need > h->part_bytes) {
            cudaFree(h->part);
            h->part = nullptr;
            h->part_bytes = 0;
            CF_CUDA(cudaMalloc(&h->part, need));
            h->part_bytes = need;
        }
        prm.part = h->part;
    }
    if (h->reserve_only) return CF_OK;
    kern<<<dim3(tiles * K, h->nseries), NT, smem, h->stream>>>(prm);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    h->fit_launches++;
    if (K > 1) {
        fit_finalize_kernel<NT, P><<<dim3(tiles, h->nseries), NT, 0, h->stream>>>(prm);
        CF_CUDA(cudaGetLastError());
        h->launches++;
    }
    return CF_OK;
}

}  // namespace

// CF_METHOD_AUTO: the tree pays off once a window holds a few hundred elements
bool method_is_tree(const cf_lowess* h)
{
    return h->method == CF_METHOD_TREE || (h->method == CF_METHOD_AUTO && h->ns >= 1024);
}

int launch_fit(cf_lowess* h, bool use_rw)
{
    if (h->n < 2 || h->end <= h->begin) return CF_OK;
    if (method_is_tree(h)) return launch_fit_tree(h, use_rw);
    FitParams prm;
    prm.x = h->x; prm.y = h->y; prm.rw = use_rw ? h->rw : nullptr; prm.ys = h->ys;
    prm.w = h->win; prm.n = h->n; prm.ns = h->ns; prm.begin = h->begin; prm.end = h->end;
    prm.ksplit = 1; prm.part = nullptr; prm.fidx = nullptr;
    if (h->delta > 0.0) { prm.fidx = h->fidx; prm.begin = h->slot_begin; prm.end = h->slot_end; }
    if (h->sms <= 0) {
        int sms = 0;
        CF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
        h->sms = sms;
    }
    switch (h->fit_variant) {
        // development variants (env CF_FIT_VARIANT); all keep T = 512.  Measured on
        // B200 at n = 1e6, ns = 5e4: every shape lands within 2 % of the default --
        // the kernel is limited by register-file operand bandwidth of the FP64 pipe
        // (DFMA with three fresh 64-bit register operands issues at 2/3 rate), not
        // by occupancy, chunk size or points per thread.
        case 1: return launch_variant<128, 4, 2, 3, 2>(h, prm, h->sms);
        case 2: return launch_variant<64, 8, 4, 4, 2>(h, prm, h->sms);
        default: return launch_variant<128, 4, 2, 4, 2>(h, prm, h->sms);
    }
}

}  // namespace cf
