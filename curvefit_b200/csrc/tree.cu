// tree.cu -- the moment-tree evaluator of one lowest() pass (SURVEY 8f row N2,
// DESIGN 3.5): an O(n log ns) replacement for the O(n * ns) sweep of
// src/curvefit_regression.f90:542-584 that produces the same fitted values to
// rounding (measured <= 3e-13 relative against the oracle; tolerance 1e-10).
//
// Idea.  Inside a window the tricube weight is a POLYNOMIAL on either side of
// xs: with s = (x_j - xs)/h,
//     w = (1 + s^3)^3  for x_j < xs,      w = (1 - s^3)^3  for x_j > xs,
// and the reference's three discrete decisions (r <= h1 -> w = 1, r > h9 ->
// w = 0, sweep exit) are the integer ranges [lo9, lo1), [lo1, hi1], (hi1, hi9]
// that windows.cu already computes bit-exactly.  Hence every sum lowest() forms,
//     S0 = sum w rw,  S1 = sum w rw s,  S2 = sum w rw s^2,
//     T0 = sum w rw y,  T1 = sum w rw s y,
// restricted to an index block that lies wholly inside one of the three
// ranges, is a fixed linear combination of the block's power moments
//     A_k = sum rw v^k (k = 0..11),   B_k = sum rw y v^k (k = 0..10),
// v = (x - c)/d the block's own centred, scaled coordinate (|v| <= 1).  The
// combination's coefficients are the Taylor coefficients of the degree-9
// polynomial about the block centre -- computed per (point, block) in ~110
// FP64 instructions, whatever the block's size.
//
// Structure.  A binary tree over leaves of 16 consecutive elements: leaf
// moments from the data, parent moments from the children's by an exact
// affine change of variable (binomial shift; |a| + |b| <= 1, so it is
// numerically benign).  Each range is cut segment-tree fashion into
// O(log ns) nodes plus < 16 elements at either end that are evaluated
// directly with the reference's formula.  Because rw changes every
// robustness iteration, the tree is rebuilt per pass (O(n); 5.4 of the 22.5 ms of
// a pass at n = 1e8), one launch per block of leaves (tree_block_kernel).
//
// Chains.  A range's two boundary paths would cost one node per binary level.  On top of
// the binary tree, chain level g = 1, 2, ... groups the units of binary level ulev[g] into
// blocks of 2^cbits[g] and gives every unit t two more nodes: PRE_g[t] = the units from its
// block's start up to t, SUF_g[t] = the units from t to its block's end, each merged into ONE
// node in its own centred coordinate (so it lies wholly inside the window whenever the range
// covers it).  A boundary path then climbs cbits[g] binary levels per node (suffix on the
// left path, prefix on the right path); only the top, where both ends fall into one block,
// uses binary nodes.  The launcher picks the geometry from ns (one level of 16 bits at ns = 5e6):
// n = 1e8, ns = 5e6: 6.5 tricube-node steps per warp instead of ~36 nodes per point on the plain
// binary tree.  The w = 1 core does not walk at all: its whole leaves come from per-block prefix sums
// of the five low-order moments (build_core_prefix).
//
// This backend is separate from fit.cu (the direct kernel that carries the
// FP64 roofline claim); cf_lowess_set_method() selects it.
#include <stdlib.h>

#include "common.cuh"

namespace cf {

constexpr int kLeaf = 16;       // elements per leaf
constexpr int kNodeLen = 32;    // doubles per node: c, d, A[12], B[11], pad
constexpr int kKA = 12, kKB = 11;
constexpr int kMaxLevels = 30;
constexpr int kMaxGroups = 4;              // chain levels

struct TreeParams {
    const double* x;
    const double* y;
    const double* rw;           // nullptr: all ones (first pass)
    double* ys;
    double* nodes;              // [series][total_nodes][kNodeLen]
    const int32_t* span;        // [series]{min lo9, max hi9} over the shard's points (windows.cu)
    const int32_t* fidx;        // delta > 0: output slot -> point (delta.cu); begin = 0, end = #fitted
    Windows w;
    int32_t n, ns, begin, end;
    int32_t nlevels;
    int32_t cnt[kMaxLevels];    // nodes per level
    int64_t off[kMaxLevels];    // first node of the level
    int64_t total;              // nodes per series (binary tree + chains)
    int32_t ngroups;            // chain levels g = 1 .. ngroups
    int32_t cbits[kMaxGroups + 1];   // chain level g groups 2^cbits[g] units into a block
    int32_t ulev[kMaxGroups + 2];    // its units are the binary nodes of level ulev[g]; ulev[g+1] = ulev[g] + cbits[g]
    int64_t coff[kMaxGroups + 1];    // coff[g]: first SUF_g node; PRE_g follows after cnt[ulev[g]] nodes
    // core prefix sums (see build_core_prefix): per leaf kQLen doubles, per block of 2^qbits leaves a
    // header (centre, half-width); qbits == 0: not built
    double* qsum;
    double* qhdr;
    int32_t qbits;
};

constexpr int kQLen = 6;             // A0, A1, A2, B0, B1, pad

__device__ __forceinline__ double2 ld2(const double* p)
{
    return __ldg(reinterpret_cast<const double2*>(p));
}

// first / last element of node (level, k)
__device__ __forceinline__ void node_span(int level, int64_t k, int32_t n, int32_t& j0, int32_t& j1)
{
    const int64_t a = (k << level) * kLeaf;
    int64_t b = ((k + 1) << level) * kLeaf;
    if (b > n) b = n;
    j0 = (int32_t)a;
    j1 = (int32_t)(b - 1);
}

// Only nodes that lie wholly inside the element span of the shard's windows can be
// reached by tree_eval_kernel's walks (they tile sub-ranges of [lo9, hi9]); a parent
// inside the span has both children inside it.  On a range-sharded series this keeps
// the per-pass rebuild proportional to the shard plus its halo, not to n.
__device__ __forceinline__ bool node_needed(const TreeParams& prm, int series, int32_t j0, int32_t j1)
{
    const int32_t lo = __ldg(prm.span + 2 * series), hi = __ldg(prm.span + 2 * series + 1);
    return j0 >= lo && j1 <= hi;
}

// ---- build: leaves -------------------------------------------------------------------
// centre and half-width of prefix block k (2^qbits leaves), as stored in its header
__device__ __forceinline__ void core_block_coord(const TreeParams& prm, const double* __restrict__ x, int32_t k,
                                                 double& cB, double& dB)
{
    const int64_t e0 = ((int64_t)k << prm.qbits) * kLeaf;
    int64_t e1 = ((int64_t)(k + 1) << prm.qbits) * kLeaf;
    if (e1 > prm.n) e1 = prm.n;
    const double xf = x[e0], xl = x[e1 - 1];
    cB = 0.5 * xf + 0.5 * xl;
    dB = 0.5 * (xl - xf);
}

__device__ __forceinline__ void build_leaf(const TreeParams& prm, int series, int32_t leaf)
{
    const size_t soff = (size_t)series * (size_t)prm.n;
    const double* __restrict__ x = prm.x + soff;
    const double* __restrict__ y = prm.y + soff;
    const double* __restrict__ rw = prm.rw ? prm.rw + soff : nullptr;
    double* __restrict__ nd = prm.nodes + ((size_t)series * prm.total + leaf) * kNodeLen;

    int32_t j0, j1;
    node_span(0, leaf, prm.n, j0, j1);
    // core prefix sums (build_core_prefix): the leaf's entry starts as its own moments A0..A2, B0,
    // B1 in the block's coordinate -- zero for a leaf the shard's windows cannot reach
    double2* qe = prm.qbits ? reinterpret_cast<double2*>(prm.qsum + ((size_t)series * prm.cnt[0] + leaf) * kQLen) : nullptr;
    if (!node_needed(prm, series, j0, j1)) {
        if (qe) { qe[0] = make_double2(0.0, 0.0); qe[1] = make_double2(0.0, 0.0); qe[2] = make_double2(0.0, 0.0); }
        return;
    }
    const double xf = x[j0], xl = x[j1];
    const double c = 0.5 * xf + 0.5 * xl;
    const double d = 0.5 * (xl - xf);
    const double inv = d > 0.0 ? 1.0 / d : 0.0;
    double A[kKA], Bm[kKB];
#pragma unroll
    for (int k = 0; k < kKA; ++k) A[k] = 0.0;
#pragma unroll
    for (int k = 0; k < kKB; ++k) Bm[k] = 0.0;
    for (int32_t j = j0; j <= j1; ++j) {
        const double v = (x[j] - c) * inv;
        double p = rw ? rw[j] : 1.0;
        double py = p * y[j];
#pragma unroll
        for (int k = 0; k < kKA; ++k) {
            A[k] += p;
            if (k < kKB) Bm[k] += py;
            p *= v;
            py *= v;
        }
    }
    double2* o = reinterpret_cast<double2*>(nd);
    o[0] = make_double2(c, d);
#pragma unroll
    for (int k = 0; k < kKA / 2; ++k) o[1 + k] = make_double2(A[2 * k], A[2 * k + 1]);
#pragma unroll
    for (int k = 0; k < 5; ++k) o[7 + k] = make_double2(Bm[2 * k], Bm[2 * k + 1]);
    o[12] = make_double2(Bm[10], 0.0);
    if (qe) {
        double cB, dB;
        core_block_coord(prm, x, leaf >> prm.qbits, cB, dB);
        const double invB = dB > 0.0 ? 1.0 / dB : 0.0;
        const double a = d * invB, b = (c - cB) * invB;
        qe[0] = make_double2(A[0], fma(b, A[0], a * A[1]));
        qe[1] = make_double2(fma(b * b, A[0], fma(2.0 * a * b, A[1], a * a * A[2])), Bm[0]);
        qe[2] = make_double2(fma(b, Bm[0], a * Bm[1]), 0.0);
    }
}

__global__ void __launch_bounds__(128)
tree_leaf_kernel(const TreeParams prm)
{
    const int32_t leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf < prm.cnt[0]) build_leaf(prm, blockIdx.y, leaf);
}

// ---- build: parents ------------------------------------------------------------------
// Moments of one child re-expressed in the parent's coordinate v_p = b + a v_c and
// added to the parent's:  M_k += sum_m C(k,m) b^(k-m) a^m M^c_m.
template <int K>
__device__ __forceinline__ void shift_add(const double* src, double a, double b,
                                          double* acc)
{
    double m[K];
    double ap = 1.0;
#pragma unroll
    for (int k = 0; k < K; ++k) {
        m[k] = src[k] * ap;
        ap *= a;
    }
    // binomial transform by repeated neighbour accumulation (Pascal's triangle)
#pragma unroll
    for (int r = 1; r < K; ++r) {
#pragma unroll
        for (int k = K - 1; k >= r; --k) m[k] = fma(b, m[k - 1], m[k]);
    }
#pragma unroll
    for (int k = 0; k < K; ++k) acc[k] += m[k];
}

__device__ __forceinline__ void build_parent(const TreeParams& prm, const double* __restrict__ x,
                                             double* base, int level, int32_t k, int series)
{
    const double* ch = base + (prm.off[level - 1] + 2 * (int64_t)k) * kNodeLen;
    double* nd = base + (prm.off[level] + k) * kNodeLen;
    int32_t j0, j1;
    node_span(level, k, prm.n, j0, j1);
    if (!node_needed(prm, series, j0, j1)) return;
    const double xf = x[j0], xl = x[j1];
    const double c = 0.5 * xf + 0.5 * xl;
    const double d = 0.5 * (xl - xf);
    const double inv = d > 0.0 ? 1.0 / d : 0.0;
    double A[kKA], Bm[kKB];
#pragma unroll
    for (int q = 0; q < kKA; ++q) A[q] = 0.0;
#pragma unroll
    for (int q = 0; q < kKB; ++q) Bm[q] = 0.0;
    const int nch = (2 * (int64_t)k + 1 < prm.cnt[level - 1]) ? 2 : 1;
    for (int s = 0; s < nch; ++s) {
        const double* cn = ch + (size_t)s * kNodeLen;   // written by this kernel in tree_top: plain loads
        const double a = cn[1] * inv;
        const double b = (cn[0] - c) * inv;
        shift_add<kKA>(cn + 2, a, b, A);
        shift_add<kKB>(cn + 2 + kKA, a, b, Bm);
    }
    double2* o = reinterpret_cast<double2*>(nd);
    o[0] = make_double2(c, d);
#pragma unroll
    for (int q = 0; q < kKA / 2; ++q) o[1 + q] = make_double2(A[2 * q], A[2 * q + 1]);
#pragma unroll
    for (int q = 0; q < 5; ++q) o[7 + q] = make_double2(Bm[2 * q], Bm[2 * q + 1]);
    o[12] = make_double2(Bm[10], 0.0);
}

__global__ void __launch_bounds__(128)
tree_merge_kernel(const TreeParams prm, int level)
{
    const int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= prm.cnt[level]) return;
    const double* __restrict__ x = prm.x + (size_t)blockIdx.y * (size_t)prm.n;
    double* base = prm.nodes + (size_t)blockIdx.y * prm.total * kNodeLen;
    build_parent(prm, x, base, level, k, blockIdx.y);
}

// all levels from `level0` up in one CTA per series (the first of them holds <= 256 nodes: one per
// thread; with wider levels in here the CTA's serial trips cost 125 us per pass at n = 1e8)
__global__ void __launch_bounds__(256)
tree_top_kernel(const TreeParams prm, int level0)
{
    const double* __restrict__ x = prm.x + (size_t)blockIdx.x * (size_t)prm.n;
    double* base = prm.nodes + (size_t)blockIdx.x * prm.total * kNodeLen;
    for (int level = level0; level < prm.nlevels; ++level) {
        for (int32_t k = threadIdx.x; k < prm.cnt[level]; k += blockDim.x)
            build_parent(prm, x, base, level, k, blockIdx.x);
        __threadfence_block();
        __syncthreads();
    }
}

// ---- build: chains -------------------------------------------------------------------
struct NodeRegs { double c, d, A[kKA], B[kKB]; };

__device__ __forceinline__ void node_load(const double* nd, NodeRegs& r)
{
    r.c = nd[0]; r.d = nd[1];
#pragma unroll
    for (int k = 0; k < kKA; ++k) r.A[k] = nd[2 + k];
#pragma unroll
    for (int k = 0; k < kKB; ++k) r.B[k] = nd[2 + kKA + k];
}
__device__ __forceinline__ void node_prefetch(const double* nd)
{
    asm volatile("prefetch.global.L1 [%0];" :: "l"(nd));
    asm volatile("prefetch.global.L1 [%0];" :: "l"(nd + 16));
}
__device__ __forceinline__ void node_store(double* nd, const NodeRegs& r)
{
    double2* o = reinterpret_cast<double2*>(nd);
    o[0] = make_double2(r.c, r.d);
#pragma unroll
    for (int q = 0; q < kKA / 2; ++q) o[1 + q] = make_double2(r.A[2 * q], r.A[2 * q + 1]);
#pragma unroll
    for (int q = 0; q < 5; ++q) o[7 + q] = make_double2(r.B[2 * q], r.B[2 * q + 1]);
    o[12] = make_double2(r.B[10], 0.0);
}
// out = the union of two disjoint neighbouring nodes, in its own centred coordinate
__device__ __forceinline__ void node_merge(const NodeRegs& p, const NodeRegs& q, NodeRegs& out)
{
    const double xf = fmin(p.c - p.d, q.c - q.d), xl = fmax(p.c + p.d, q.c + q.d);
    out.c = 0.5 * xf + 0.5 * xl;
    out.d = 0.5 * (xl - xf);
    const double inv = out.d > 0.0 ? 1.0 / out.d : 0.0;
#pragma unroll
    for (int k = 0; k < kKA; ++k) out.A[k] = 0.0;
#pragma unroll
    for (int k = 0; k < kKB; ++k) out.B[k] = 0.0;
    shift_add<kKA>(p.A, p.d * inv, (p.c - out.c) * inv, out.A);
    shift_add<kKB>(p.B, p.d * inv, (p.c - out.c) * inv, out.B);
    shift_add<kKA>(q.A, q.d * inv, (q.c - out.c) * inv, out.A);
    shift_add<kKB>(q.B, q.d * inv, (q.c - out.c) * inv, out.B);
}

// ---- evaluation ------------------------------------------------------------------------
// Segment-tree walk over the leaf range [la, lb): yields, one node per call, the
// O(log) maximal aligned nodes that tile it (bottom-up, the classic iterative
// decomposition).  Per-thread state only; no lists in memory.
struct Walk {
    int32_t la, lb;
    int level, sub;
    __device__ __forceinline__ void start(int32_t a, int32_t b) { la = a; lb = b; level = 0; sub = 0; }
    __device__ __forceinline__ void start_at(int32_t a, int32_t b, int lv0) { la = a; lb = b; level = lv0; sub = 0; }
    __device__ __forceinline__ bool next(int& lv, int32_t& idx)
    {
        while (la < lb) {
            if (sub == 0) {
                sub = 1;
                if (la & 1) { lv = level; idx = la++; return true; }
            }
            sub = 0;
            const bool t = lb & 1;
            const int32_t id = lb - 1;
            const int l0 = level;
            if (t) lb = id;
            la >>= 1; lb >>= 1; ++level;
            if (t) { lv = l0; idx = id; return true; }
        }
        return false;
    }
};

// Small levels (too few units to fill the GPU with segments): one thread per (unit t,
// direction), PRE_g[t] / SUF_g[t] merged from the few aligned binary nodes that tile it.
__global__ void __launch_bounds__(128)
tree_chain_entry_kernel(const TreeParams prm, int g)
{
    const int ulev = prm.ulev[g];
    const int32_t cm = (1 << prm.cbits[g]) - 1;
    const int32_t ucnt = prm.cnt[ulev];
    const int32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    const int32_t t = tid >> 1;
    const int dir = tid & 1;
    if (t >= ucnt) return;
    const int series = blockIdx.y;
    const int32_t t0 = t & ~cm, t1 = min(t0 + cm + 1, ucnt);
    const int32_t ra = dir == 0 ? t0 : t, rb = dir == 0 ? t + 1 : t1;     // units [ra, rb)
    {   // only what the windows of this shard can reach
        int32_t j0, j1, k0, k1;
        node_span(ulev, ra, prm.n, j0, k0);
        node_span(ulev, rb - 1, prm.n, k1, j1);
        if (!node_needed(prm, series, j0, j1)) return;
    }
    double* base = prm.nodes + (size_t)series * prm.total * kNodeLen;
    double* out = base + (prm.coff[g] + (dir == 0 ? (int64_t)ucnt : 0) + t) * kNodeLen;
    Walk w;
    w.start_at(ra, rb, ulev);
    int lv;
    int32_t idx;
    NodeRegs acc, u, nx;
    w.next(lv, idx);
    node_load(base + (prm.off[lv] + idx) * kNodeLen, acc);
    while (w.next(lv, idx)) {
        node_load(base + (prm.off[lv] + idx) * kNodeLen, u);
        node_merge(acc, u, nx);
        acc = nx;
    }
    node_store(out, acc);
}

// One thread per (segment of <= 16 units, direction).  PRE: the units from the block's start
// up to the segment come from the binary tree (<= a few aligned nodes), then the segment is
// swept left to right, one merge and one stored node per unit; SUF mirrors it.
constexpr int kChainSeg = 16;

__device__ __forceinline__ void chain_segment(const TreeParams& prm, int g, int series, int32_t seg,
                                              int32_t sg, int dir)
{
    const int ulev = prm.ulev[g];
    const int32_t cm = (1 << prm.cbits[g]) - 1;
    const int32_t ucnt = prm.cnt[ulev];
    const int32_t s0 = sg * seg, s1 = min(s0 + seg, ucnt);                  // units [s0, s1)
    const int32_t t0 = s0 & ~cm, t1 = min(t0 + cm + 1, ucnt);               // their block
    {   // only what the windows of this shard can reach
        int32_t j0, j1, k0, k1;
        node_span(ulev, s0, prm.n, j0, k0);
        node_span(ulev, s1 - 1, prm.n, k1, j1);
        const int32_t lo = __ldg(prm.span + 2 * series), hi = __ldg(prm.span + 2 * series + 1);
        if (j1 < lo || j0 > hi) return;
    }
    double* base = prm.nodes + (size_t)series * prm.total * kNodeLen;
    const double* units = base + prm.off[ulev] * kNodeLen;
    double* suf = base + prm.coff[g] * kNodeLen;
    double* pre = suf + (size_t)ucnt * kNodeLen;
    NodeRegs acc, u, nx;
    bool have = false;
    {   // what lies between the block's edge and the segment
        Walk w;
        if (dir == 0) w.start_at(t0, s0, ulev); else w.start_at(s1, t1, ulev);
        int lv;
        int32_t idx;
        while (w.next(lv, idx)) {
            node_load(base + (prm.off[lv] + idx) * kNodeLen, u);
            if (have) { node_merge(acc, u, nx); acc = nx; } else { acc = u; have = true; }
        }
    }
    if (dir == 0) {
        for (int32_t t = s0; t < s1; ++t) {
            node_load(units + (size_t)t * kNodeLen, u);
            if (t + 1 < s1) node_prefetch(units + (size_t)(t + 1) * kNodeLen);
            if (have) { node_merge(acc, u, nx); acc = nx; } else { acc = u; have = true; }
            node_store(pre + (size_t)t * kNodeLen, acc);
        }
    } else {
        for (int32_t t = s1 - 1; t >= s0; --t) {
            node_load(units + (size_t)t * kNodeLen, u);
            if (t > s0) node_prefetch(units + (size_t)(t - 1) * kNodeLen);
            if (have) { node_merge(acc, u, nx); acc = nx; } else { acc = u; have = true; }
            node_store(suf + (size_t)t * kNodeLen, acc);
        }
    }
}

__global__ void __launch_bounds__(128)
tree_chain_kernel(const TreeParams prm, int g)
{
    const int32_t seg = min(kChainSeg, 1 << prm.cbits[g]);                   // never across blocks
    const int32_t nseg = (prm.cnt[prm.ulev[g]] + seg - 1) / seg;
    const int32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    if ((tid >> 1) < nseg) chain_segment(prm, g, blockIdx.y, seg, tid >> 1, tid & 1);
}

// ---- build: core prefix sums ------------------------------------------------------------------
// The w = 1 core [lo1, hi1] (:547-551) holds ~0.1 % of the window but its whole leaves are a range of
// their own: at ns = 5e6 a dozen binary nodes per point, each a step of the walk -- 19 % of the
// evaluator's time for a sum that only needs the moments A0..A2, B0, B1.  Per block of 2^qbits
// leaves (the block of tree_block_kernel) and leaf t the inclusive prefix Q[t] of those five
// moments over the block's leaves up to t is kept in the BLOCK's centred coordinate; the core's
// whole leaves inside one block are then Q[t1 - 1] - Q[t0 - 1]: two 48-byte loads instead of a walk.
// build_leaf leaves every leaf's own five moments (block coordinate; zeros for leaves the shard's
// windows cannot reach) in its entry; the first warp of the block's CTA then scans them in place.
__device__ __forceinline__ void build_core_prefix(const TreeParams& prm, int series, int32_t blk0,
                                                  int32_t blk1, int cb)
{
    constexpr unsigned kFull = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int per = (1 << cb) >> 5;                               // leaves per lane (cb >= 5)
    const int32_t first = blk0 + lane * per;
    double* q = prm.qsum + ((size_t)series * prm.cnt[0] + first) * kQLen;   // written by build_leaf: plain loads
    double tot[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    for (int r = 0; r < per && first + r < blk1; ++r) {
#pragma unroll
        for (int k = 0; k < 5; ++k) tot[k] += q[(size_t)r * kQLen + k];
    }
    double run[5];                                                // exclusive scan of the lane totals
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        double v = tot[k];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double u = __shfl_up_sync(kFull, v, o);
            if (lane >= o) v += u;
        }
        run[k] = v - tot[k];
    }
    for (int r = 0; r < per && first + r < blk1; ++r) {
        double2* o = reinterpret_cast<double2*>(q + (size_t)r * kQLen);
        const double2 e0 = o[0], e1 = o[1], e2 = o[2];
        run[0] += e0.x; run[1] += e0.y; run[2] += e1.x; run[3] += e1.y; run[4] += e2.x;
        o[0] = make_double2(run[0], run[1]);
        o[1] = make_double2(run[2], run[3]);
        o[2] = make_double2(run[4], 0.0);
    }
    if (lane == 0) {
        double cB, dB;
        core_block_coord(prm, prm.x + (size_t)series * (size_t)prm.n, blk0 >> cb, cB, dB);
        double2* hd = reinterpret_cast<double2*>(prm.qhdr + ((size_t)series * ((prm.cnt[0] >> cb) + 1) + (blk0 >> cb)) * 2);
        *hd = make_double2(cB, dB);
    }
}

// ---- build: one launch per block of leaves ------------------------------------------------
// One CTA per block of 2^cb leaves builds everything that lives inside the block: its
// leaves, the binary levels 1 .. cb above them and (chain != 0) the level-1 chain PRE_1 /
// SUF_1, phase after phase with the block's nodes still in L2 -- the separate kernels
// re-read each level from HBM (leaf + merges + chain: 15.2 GB per pass at n = 1e8, fused:
// 8.8 GB, of which 6.4 GB are the stores).
// What it buys is launches, not bandwidth: at n = 1e6 a smooth drops from 2.35 to 1.71 ms
// (79 -> 59 launches), at n = 1e8 the build stays at ~5.3 ms per pass.  ncu there: l1tex
// throughput 87 %, DRAM 23 % -- a node per thread is 25 loads / 13 stores at a stride of
// 256 B between lanes, 32 lines per warp instruction.  Moving the nodes through a per-warp
// shared-memory stage (the warp fetching lane i's node with 25 lanes at once) was measured
// and dropped: staging loads and stores 147 ms per smooth instead of 120 (the 32-step
// fetch loop serialises the load latency), staging only the stores and the leaf inputs
// 120.0-123.3 ms, no gain.
// NT = 32: one warp per block, __syncwarp between the phases -- no warp waits at a barrier
// for the few lanes that build the upper levels, and 16 such CTAs fit an SM; best for many
// blocks.  NT = 128: shortest critical path per block; best when the blocks are few.
template <int NT>
__global__ void __launch_bounds__(NT, 512 / NT)
tree_block_kernel(const TreeParams prm, int cb, int chain, int seg)
{
    const int series = blockIdx.y;
    const int32_t blk0 = blockIdx.x << cb;
    const int32_t blk1 = min(blk0 + (1 << cb), prm.cnt[0]);
    {   // a block none of whose elements the shard's windows reach has nothing to build
        const int32_t lo = __ldg(prm.span + 2 * series), hi = __ldg(prm.span + 2 * series + 1);
        int32_t e0, e1, k0, k1;
        node_span(0, blk0, prm.n, e0, k0);
        node_span(0, blk1 - 1, prm.n, k1, e1);
        if (e1 < lo || e0 > hi) return;
    }
    auto sync = [] {
        __threadfence_block();
        if (NT == 32) __syncwarp(); else __syncthreads();
    };
    for (int32_t leaf = blk0 + threadIdx.x; leaf < blk1; leaf += NT) build_leaf(prm, series, leaf);
    sync();
    if (prm.qbits && threadIdx.x < 32) build_core_prefix(prm, series, blk0, blk1, cb);
    const double* __restrict__ x = prm.x + (size_t)series * (size_t)prm.n;
    double* base = prm.nodes + (size_t)series * prm.total * kNodeLen;
    for (int level = 1; level <= cb && level < prm.nlevels; ++level) {
        const int32_t k0 = blk0 >> level, k1 = ((blk1 - 1) >> level) + 1;
        for (int32_t k = k0 + threadIdx.x; k < k1; k += NT) build_parent(prm, x, base, level, k, series);
        sync();
    }
    if (chain) {
        const int32_t sg0 = blk0 / seg, sg1 = (blk1 - 1) / seg + 1;
        for (int32_t task = threadIdx.x; task < 2 * (sg1 - sg0); task += NT)
            chain_segment(prm, 1, series, seg, sg0 + (task >> 1), task & 1);
    }
}

// The nodes that tile the leaf range [la, lb): per chain level at most one suffix node (left
// end) and one prefix node (right end), then the binary walk over what is left where both
// ends share a block.  Yields node indices into TreeParams::nodes (total < n, so 32 bits).
// Five registers of state: the evaluator's occupancy is bought with them.
struct Chain {
    int32_t a, b;   // what is left to tile, in units of the level being climbed
    int g;          // chain level 1..ngroups while climbing the chains, 0 = binary tail
    int st;         // chains: 0 test / left end, 1 right end, 2 climb;  tail: 0 / 1 as Walk::sub
    int lv;         // binary level of a, b in the tail
    __device__ __forceinline__ void start(int32_t la, int32_t lb)
    {
        a = la; b = lb; g = 1; st = 0; lv = 0;
        if (a >= b) { a = b = 0; g = 0; }
    }
    __device__ __forceinline__ bool next(const TreeParams& prm, int32_t& node)
    {
        while (g) {
            const int cb = g <= prm.ngroups ? prm.cbits[g] : 0;
            const int32_t cm = (1 << cb) - 1;
            if (st == 0) {
                if (cb == 0 || (a >> cb) == ((b - 1) >> cb)) {
                    lv = prm.ulev[g];
                    g = 0;
                    break;
                }
                st = 1;
                if (a & cm) { node = (int32_t)prm.coff[g] + a; a = (a | cm) + 1; return true; }
            }
            if (st == 1) {
                st = 2;
                if (b & cm) { node = (int32_t)prm.coff[g] + prm.cnt[prm.ulev[g]] + (b - 1); b &= ~cm; return true; }
            }
            a >>= cb; b >>= cb; ++g; st = 0;
            if (a >= b) { a = b = 0; g = 0; return false; }
        }
        while (a < b) {                                   // as Walk::next
            if (st == 0) {
                st = 1;
                if (a & 1) { node = (int32_t)prm.off[lv] + a; ++a; return true; }
            }
            st = 0;
            const bool t = b & 1;
            const int32_t id = b - 1;
            const int l0 = lv;
            if (t) b = id;
            a >>= 1; b >>= 1; ++lv;
            if (t) { node = (int32_t)prm.off[l0] + id; return true; }
        }
        return false;
    }
};

// [a, b) -> direct head [a, hd), whole leaves [la, lb), direct tail [tl, b)
__device__ __forceinline__ void split_range(int32_t a, int32_t b, int32_t& hd, int32_t& tl,
                                            int32_t& la, int32_t& lb)
{
    const int32_t up = (a + kLeaf - 1) / kLeaf * kLeaf;
    const int32_t dn = b / kLeaf * kLeaf;
    hd = min(b, up);
    tl = max(hd, dn);
    la = up / kLeaf;
    lb = dn / kLeaf;
    if (la >= lb) { la = lb = 0; }
}

struct Moments { double S0, S1, S2, T0, T1; };

// One node inside a tricube range.  nsig = +1 left of xs (w = (1 + s^3)^3),
// -1 right of xs (w = (1 - s^3)^3).
template <bool SH>
__device__ __forceinline__ double2 ldn(const double* p)
{
    if (SH) return *reinterpret_cast<const double2*>(p);      // staged node in shared memory
    return ld2(p);
}

template <bool SH = false>
__device__ __forceinline__ void take_tricube(const double* __restrict__ nd, double xs, double invh,
                                             double nsig, Moments& m)
{
    const double2 cd = ldn<SH>(nd);
    const double beta = (cd.x - xs) * invh;       // s at the node centre
    const double alpha = cd.y * invh;             // ds/dv
    const double b2 = beta * beta, a2 = alpha * alpha;
    // 1 + nsig (beta + alpha v)^3 = c0 + c1 v + c2 v^2 + c3 v^3
    const double c0 = fma(nsig, b2 * beta, 1.0);
    const double c1 = nsig * (3.0 * b2 * alpha);
    const double c2 = nsig * (3.0 * beta * a2);
    const double c3 = nsig * (a2 * alpha);
    // square (degree 6)
    const double d0 = c0 * c0;
    const double d1 = 2.0 * c0 * c1;
    const double d2 = fma(c1, c1, 2.0 * c0 * c2);
    const double d3 = 2.0 * fma(c0, c3, c1 * c2);
    const double d4 = fma(c2, c2, 2.0 * c1 * c3);
    const double d5 = 2.0 * c2 * c3;
    const double d6 = c3 * c3;
    // cube (degree 9): f = d * c
    double f[10];
    f[0] = d0 * c0;
    f[1] = fma(d1, c0, d0 * c1);
    f[2] = fma(d2, c0, fma(d1, c1, d0 * c2));
    f[3] = fma(d3, c0, fma(d2, c1, fma(d1, c2, d0 * c3)));
    f[4] = fma(d4, c0, fma(d3, c1, fma(d2, c2, d1 * c3)));
    f[5] = fma(d5, c0, fma(d4, c1, fma(d3, c2, d2 * c3)));
    f[6] = fma(d6, c0, fma(d5, c1, fma(d4, c2, d3 * c3)));
    f[7] = fma(d6, c1, fma(d5, c2, d4 * c3));
    f[8] = fma(d6, c2, d5 * c3);
    f[9] = d6 * c3;
    double A[kKA], Bm[kKA];
#pragma unroll
    for (int k = 0; k < kKA / 2; ++k) {
        const double2 t = ldn<SH>(nd + 2 + 2 * k);
        A[2 * k] = t.x; A[2 * k + 1] = t.y;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        const double2 t = ldn<SH>(nd + 2 + kKA + 2 * k);
        Bm[2 * k] = t.x; Bm[2 * k + 1] = t.y;     // Bm[11] is the node's zero pad
    }
    double D0 = 0.0, D1 = 0.0, D2 = 0.0, E0 = 0.0, E1 = 0.0;
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        D0 = fma(f[k], A[k], D0);
        D1 = fma(f[k], A[k + 1], D1);
        D2 = fma(f[k], A[k + 2], D2);
        E0 = fma(f[k], Bm[k], E0);
        E1 = fma(f[k], Bm[k + 1], E1);
    }
    // s = beta + alpha v:  sum w rw s = beta D0 + alpha D1, etc.
    m.S0 += D0;
    m.S1 += fma(beta, D0, alpha * D1);
    m.S2 += fma(b2, D0, fma(2.0 * alpha * beta, D1, a2 * D2));
    m.T0 += E0;
    m.T1 += fma(beta, E0, alpha * E1);
}

// One node inside the w = 1 core.
template <bool SH = false>
__device__ __forceinline__ void take_core(const double* __restrict__ nd, double xs, double invh,
                                          Moments& m)
{
    const double2 cd = ldn<SH>(nd);
    const double beta = (cd.x - xs) * invh, alpha = cd.y * invh;
    const double2 a01 = ldn<SH>(nd + 2);
    const double2 a23 = ldn<SH>(nd + 4);
    const double2 b01 = ldn<SH>(nd + 2 + kKA);
    m.S0 += a01.x;
    m.S1 += fma(beta, a01.x, alpha * a01.y);
    m.S2 += fma(beta * beta, a01.x, fma(2.0 * alpha * beta, a01.y, alpha * alpha * a23.x));
    m.T0 += b01.x;
    m.T1 += fma(beta, b01.x, alpha * b01.y);
}

// ---- warp-cooperative node staging ----------------------------------------------------------
// The 32 lanes of a warp are 32 neighbouring output points: their node lists are almost the
// same (the windows shift by about one element per point, a node covers >= 16), and the nodes
// they need at the same step of the walk are NEIGHBOURS in the node array (SUF_g[a], SUF_g[a+1]
// ...).  Instead of every lane fetching its own 256-byte node (13 LDG.128 per lane and node,
// 32x redundant L1 traffic), the warp copies the kStage consecutive nodes starting at the
// smallest index any lane asks for into shared memory with two 16-byte cp.async per lane
// (1 KB, fully coalesced), and every lane whose node is among them evaluates it from there
// (broadcast LDS).  Lanes whose node lies elsewhere wait for a later step; a lane consumes its
// own list in its own order whatever the other lanes do, so a point's sums do not depend on
// which warp it is in.  The copy for the NEXT step is in flight while this step's ~110 FP64
// instructions per node run (double buffer; the next indices are known before the evaluation).
constexpr int kStage = 4;            // nodes staged per warp and step
constexpr int kStagePad = 34;        // doubles per staged node: 272 B, so that lanes on different
                                     // slots read different banks
constexpr int kStageLen = kStage * kStagePad;
constexpr unsigned kNoNode = 0xffffffffu;

__device__ __forceinline__ void cp_async16(double* smem, const double* gmem)
{
    const unsigned sa = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// nodes [m, m + kStage) -> buf (the node array is padded by kStage nodes, so this never leaves it)
__device__ __forceinline__ void stage_nodes(double* buf, const double* __restrict__ nodes, unsigned m)
{
    const int lane = threadIdx.x & 31;
    const double* src = nodes + (size_t)m * kNodeLen;
#pragma unroll
    for (int h = 0; h < (kStage * kNodeLen) / 64; ++h) {
        const int c = lane + 32 * h;                  // 16-byte chunk of the kStage * 256 bytes
        const int q = c >> 4, o = (c & 15) * 2;
        cp_async16(buf + q * kStagePad + o, src + q * kNodeLen + o);
    }
    cp_async_commit();
}

// The walk of one phase for a whole warp.  have/node: this lane's current item; advance():
// moves the lane to its next item, returns false when its list is exhausted; tag(): per-item
// value read BEFORE advance (the side of the tricube); take(staged node, tag).
template <class Adv, class Tag, class Take>
__device__ __forceinline__ void staged_walk(const double* __restrict__ nodes, double* stage, bool& have,
                                            int32_t& node, Adv&& advance, Tag&& tag, Take&& take)
{
    constexpr unsigned kFull = 0xffffffffu;
    unsigned m = __reduce_min_sync(kFull, have ? (unsigned)node : kNoNode);
    if (m == kNoNode) return;
    int buf = 0;
    stage_nodes(stage, nodes, m);
    for (;;) {
        const unsigned cur = have ? (unsigned)node : kNoNode;
        const bool served = cur - m < (unsigned)kStage;        // cur >= m for every lane that has an item
        const double tg = tag();
        if (served) have = advance();
        const unsigned m2 = __reduce_min_sync(kFull, have ? (unsigned)node : kNoNode);
        cp_async_wait_all();
        __syncwarp();                                          // this step's nodes have landed for all lanes
        if (m2 != kNoNode) stage_nodes(stage + (buf ^ 1) * kStageLen, nodes, m2);
        if (served) take(stage + buf * kStageLen + (cur - m) * kStagePad, tg);
        if (m2 == kNoNode) break;
        buf ^= 1;
        m = m2;
    }
    __syncwarp();
}

// One thread per output point.  Three phases, each a warp-convergent loop in
// which every lane works on ITS OWN k-th item (node or element), so no lane
// evaluates an item it does not need:
//   1. tricube nodes of [lo9, lo1) and (hi1, hi9]   (~110 FP64 instr each)
//   2. core nodes of [lo1, hi1]                     (~15)
//   3. the < 16 elements at either end of each range, reference formula (~13 each)
// (Tried and dropped, B200, n = 1e8: loading one node ahead into registers, splitting the
// phases into two kernels, advancing the left and right walks together -- all within 2 % or
// slower; occupancy is what pays, see the register diet below.)
template <int NT, int MINB, int PF>
__global__ void __launch_bounds__(NT, MINB)
tree_eval_kernel(const TreeParams prm)
{
    // Register diet: the kernel waits on node loads, so occupancy is what pays (6 CTAs of
    // 128 threads per SM instead of 5: -7 %; 7: another -1 %).  Nothing but the moments, xs,
    // 1/h, the walk and the point's index lives through phase 1: the series offset is
    // rebuilt from %ctaid.y (soff(), opaque to the compiler so that it is not kept), the
    // range bounds are re-read from the window arrays at the start of each phase (ldv(),
    // volatile for the same reason) and the bandwidth h is rebuilt at the end.
    auto soff = [&]() -> size_t {
        unsigned s;
        asm volatile("mov.u32 %0, %%ctaid.y;" : "=r"(s));
        return (size_t)s * (size_t)prm.n;
    };
    auto ldv = [](const int32_t* p) -> int32_t {
        int32_t v;
        asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
        return v;
    };
    const double* __restrict__ nodes = prm.nodes + (size_t)blockIdx.y * prm.total * kNodeLen;
    // PF >= 2: per-warp double buffer of staged nodes (see staged_walk); PF == 3: only the tricube
    // nodes are staged, the core nodes (4 of their 13 16-byte pieces are needed) are read directly
    __shared__ __align__(16) double stage_all[PF >= 2 ? (NT / 32) * 2 * kStageLen : 2];
    double* stage = stage_all + (PF >= 2 ? (threadIdx.x >> 5) * 2 * kStageLen : 0);

    // tie followers copy their leader's value later (robust.cu); a shard that starts
    // inside a tie group fits its first point in the leader's place (as fit.cu)
    int32_t i;                                   // the point, -1 = this lane fits nothing
    double xs = 0.0, invh = 0.0;
    {
        const size_t so = soff();
        const int32_t slot = prm.begin + blockIdx.x * NT + threadIdx.x;
        bool valid = slot < prm.end;
        i = (valid && prm.fidx) ? __ldg(prm.fidx + slot) : slot;
        if (valid && !prm.fidx) {
            const int32_t l = prm.w.leader[so + i];
            valid = (l == i) || (i == prm.begin && l < i);
        }
        if (valid) {
            const double* __restrict__ x = prm.x + so;
            const int32_t nl = prm.w.nleft[so + i];
            xs = x[i];
            const double hl = xs - x[nl], hr = x[nl + prm.ns - 1] - xs;    // :536
            const double h = hl > hr ? hl : hr;
            invh = h > 0.0 ? 1.0 / h : 0.0;
        } else {
            i = -1;
        }
    }
    // ranges: L = [l9, l1), C = [l1, r1 + 1), R = [r1 + 1, r9 + 1)

    Moments m = {0.0, 0.0, 0.0, 0.0, 0.0};
    constexpr unsigned kFull = 0xffffffffu;

    // ---- phase 1: tricube nodes, left range then right range
    {
        Chain w;
        {
            int32_t hd, tl, la = 0, lb = 0;
            if (i >= 0) {
                const size_t pi = soff() + i;
                split_range(ldv(prm.w.lo9 + pi), ldv(prm.w.lo1 + pi), hd, tl, la, lb);
            }
            w.start(la, lb);
        }
        int phase = 0;
        int32_t node = 0;
        double nsig = 1.0;
        auto advance = [&]() -> bool {
            if (w.next(prm, node)) return true;
            if (phase == 0) {
                phase = 1;
                nsig = -1.0;
                int32_t hd, tl, la = 0, lb = 0;
                if (i >= 0) {
                    const size_t pi = soff() + i;
                    split_range(ldv(prm.w.hi1 + pi) + 1, ldv(prm.w.hi9 + pi) + 1, hd, tl, la, lb);
                }
                w.start(la, lb);
                return w.next(prm, node);
            }
            return false;
        };
        bool have = advance();
        if (PF >= 2) {
            staged_walk(nodes, stage, have, node, advance, [&]() { return nsig; },
                        [&](const double* nd, double sg) { take_tricube<true>(nd, xs, invh, sg, m); });
        } else if (PF == 0) {
            while (__any_sync(kFull, have)) {
                if (have) {
                    take_tricube(nodes + (size_t)node * kNodeLen, xs, invh, nsig, m);
                    have = advance();
                }
            }
        } else {
            // the index of the node after this one is known before this one is evaluated:
            // ask for its two lines while the ~110 FP64 instructions of this one run
            while (__any_sync(kFull, have)) {
                if (have) {
                    const double* cur = nodes + (size_t)node * kNodeLen;
                    const double sg = nsig;
                    have = advance();
                    if (have) {
                        const double* nx = nodes + (size_t)node * kNodeLen;
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(nx));
                        asm volatile("prefetch.global.L1 [%0];" :: "l"(nx + 16));
                    }
                    take_tricube(cur, xs, invh, sg, m);
                }
            }
        }
    }
    // ---- phase 2: core nodes
    if (PF != 4) {
        Chain w;
        {
            int32_t hd, tl, la = 0, lb = 0;
            if (i >= 0) {
                const size_t pi = soff() + i;
                split_range(ldv(prm.w.lo1 + pi), ldv(prm.w.hi1 + pi) + 1, hd, tl, la, lb);
            }
            // whole leaves that lie in at most three prefix blocks: differences of the blocks'
            // core prefix sums (build_core_prefix) instead of a walk over a dozen nodes
            const int q = prm.qbits;
            if (q && la < lb && ((lb - 1) >> q) - (la >> q) <= 2) {
                unsigned series;
                asm volatile("mov.u32 %0, %%ctaid.y;" : "=r"(series));
                const double* __restrict__ qs = prm.qsum + (size_t)series * prm.cnt[0] * kQLen;
                const double* __restrict__ qh = prm.qhdr + (size_t)series * ((prm.cnt[0] >> q) + 1) * 2;
                for (int32_t k = la >> q; k <= (lb - 1) >> q; ++k) {
                    const int32_t t0 = max(la, k << q), t1 = min(lb, (k + 1) << q);
                    const double2 cd = ld2(qh + 2 * (size_t)k);
                    const double* hi = qs + (size_t)(t1 - 1) * kQLen;
                    double2 a01 = ld2(hi), a2b0 = ld2(hi + 2), b1 = ld2(hi + 4);
                    if (t0 > (k << q)) {
                        const double* lo = qs + (size_t)(t0 - 1) * kQLen;
                        const double2 l01 = ld2(lo), l2 = ld2(lo + 2), l4 = ld2(lo + 4);
                        a01.x -= l01.x; a01.y -= l01.y; a2b0.x -= l2.x; a2b0.y -= l2.y; b1.x -= l4.x;
                    }
                    const double beta = (cd.x - xs) * invh, alpha = cd.y * invh;
                    m.S0 += a01.x;
                    m.S1 += fma(beta, a01.x, alpha * a01.y);
                    m.S2 += fma(beta * beta, a01.x, fma(2.0 * alpha * beta, a01.y, alpha * alpha * a2b0.x));
                    m.T0 += a2b0.y;
                    m.T1 += fma(beta, a2b0.y, alpha * b1.x);
                }
                la = lb = 0;
            }
            w.start(la, lb);
        }
        int32_t node = 0;
        bool have = w.next(prm, node);
        if (PF == 2) {
            staged_walk(nodes, stage, have, node, [&]() { return w.next(prm, node); }, []() { return 0.0; },
                        [&](const double* nd, double) { take_core<true>(nd, xs, invh, m); });
        } else {
            while (__any_sync(kFull, have)) {
                if (have) {
                    take_core(nodes + (size_t)node * kNodeLen, xs, invh, m);
                    have = w.next(prm, node);
                }
            }
        }
    }
    // ---- phase 3: range ends, element by element (:542-556 per pair)
    if (PF != 5) {
        // adjacent pieces are merged: [l9,hdL) | [tlL,l1)+[l1,hdC) | [tlC,r1+1)+[r1+1,hdR) | [tlR,r9+1).
        // Two independent streams per lane -- A: pieces 1, 2 (left of the core and its head),
        // B: pieces 3, 4 -- with their own accumulators, walked by ONE counter k: element k of a
        // stream is j1 + k in its first piece, j2 + (k - c1) in its second.  Inside A an element
        // belongs to the w = 1 core iff j >= l1, inside B iff j <= r1 (:547-551).
        const bool valid = i >= 0;
        // the three arrays as per-thread pointers the compiler cannot fold back into
        // (index + series offset) arithmetic: one IMAD.WIDE per load instead of seven instructions
        const double* x = prm.x + soff();
        const double* y = prm.y + soff();
        const double* rw = prm.rw ? prm.rw + soff() : nullptr;
        asm volatile("" : "+l"(x), "+l"(y), "+l"(rw));
        int32_t l9 = 0, r9 = -1, l1 = 0, r1 = -1;
        if (valid) {
            const size_t pi = soff() + i;
            l9 = ldv(prm.w.lo9 + pi); r9 = ldv(prm.w.hi9 + pi);
            l1 = ldv(prm.w.lo1 + pi); r1 = ldv(prm.w.hi1 + pi);
        }
        int32_t hdL, tlL, hdC, tlC, hdR, tlR, la, lb;
        split_range(l9, l1, hdL, tlL, la, lb);
        split_range(l1, r1 + 1, hdC, tlC, la, lb);
        split_range(r1 + 1, r9 + 1, hdR, tlR, la, lb);
        const int32_t c1a = valid ? max(hdL - l9, 0) : 0, ca = c1a + (valid ? max(hdC - tlL, 0) : 0);
        const int32_t c1b = valid ? max(hdR - tlC, 0) : 0, cb = c1b + (valid ? max(r9 + 1 - tlR, 0) : 0);
        const int32_t d2a = tlL - c1a, d2b = tlR - c1b;           // j = d2 + k in the second piece
        // last element of each stream (any readable index when the stream is empty): lanes that
        // have run out keep reading it with weight zero
        const int32_t safe = valid ? i : 0;
        const int32_t enda1 = ca > c1a ? d2a + ca - 1 : (ca > 0 ? l9 + ca - 1 : safe);
        const int32_t endb1 = cb > c1b ? d2b + cb - 1 : (cb > 0 ? tlC + cb - 1 : safe);
        Moments mb = {0.0, 0.0, 0.0, 0.0, 0.0};
        const int32_t kmax = max(ca, cb);
        for (int32_t k = 0; __any_sync(kFull, k < kmax); ++k) {
            const int32_t ja = min((k < c1a ? l9 : d2a) + k, enda1);
            const int32_t jb = min((k < c1b ? tlC : d2b) + k, endb1);
            {
                const double q = (x[ja] - xs) * invh;
                const double tt = fma(-(q * q), fabs(q), 1.0);
                double wgt = (tt * tt) * tt;
                wgt = ja >= l1 ? 1.0 : wgt;
                const double wr = rw ? wgt * rw[ja] : wgt;
                const double wq = wr * q;
                const double yj = y[ja];
                if (k < ca) {                 // predicated accumulation: no select on the weight
                    m.S0 += wr;
                    m.S1 += wq;
                    m.S2 = fma(wq, q, m.S2);
                    m.T0 = fma(wr, yj, m.T0);
                    m.T1 = fma(wq, yj, m.T1);
                }
            }
            {
                const double q = (x[jb] - xs) * invh;
                const double tt = fma(-(q * q), fabs(q), 1.0);
                double wgt = (tt * tt) * tt;
                wgt = jb <= r1 ? 1.0 : wgt;
                const double wr = rw ? wgt * rw[jb] : wgt;
                const double wq = wr * q;
                const double yj = y[jb];
                if (k < cb) {
                    mb.S0 += wr;
                    mb.S1 += wq;
                    mb.S2 = fma(wq, q, mb.S2);
                    mb.T0 = fma(wr, yj, mb.T0);
                    mb.T1 = fma(wq, yj, mb.T1);
                }
            }
        }
        m.S0 += mb.S0; m.S1 += mb.S1; m.S2 += mb.S2; m.T0 += mb.T0; m.T1 += mb.T1;
    }
    if (i >= 0) {
        const size_t so = soff();
        const double* __restrict__ x = prm.x + so;
        const int32_t nl = ldv(prm.w.nleft + so + i);
        const double hl = xs - x[nl], hr = x[nl + prm.ns - 1] - xs;        // :536, as above
        const double h = hl > hr ? hl : hr;
        const double range = x[prm.n - 1] - x[0];                          // :535
        const double y = prm.y[so + i];
        prm.ys[so + i] = fit_value(m.S0, m.S1, m.S2, m.T0, m.T1, y, h, h, 0.0, range);
    }
}

// ---- host side ---------------------------------------------------------------------------
int launch_fit_tree(cf_lowess* h, bool use_rw)
{
    if (h->n < 2 || h->end <= h->begin) return CF_OK;
    TreeParams prm = {};
    prm.x = h->x; prm.y = h->y; prm.rw = use_rw ? h->rw : nullptr; prm.ys = h->ys;
    prm.w = h->win; prm.n = h->n; prm.ns = h->ns; prm.begin = h->begin; prm.end = h->end;
    prm.fidx = nullptr;
    if (h->delta > 0.0) { prm.fidx = h->fidx; prm.begin = h->slot_begin; prm.end = h->slot_end; }
    int L = 0;
    int64_t total = 0;
    int32_t cnt = (h->n + kLeaf - 1) / kLeaf;
    for (;;) {
        prm.cnt[L] = cnt;
        prm.off[L] = total;
        total += cnt;
        ++L;
        if (cnt <= 1 || L >= kMaxLevels) break;
        cnt = (cnt + 1) / 2;
    }
    prm.nlevels = L;
    // chains: level g groups the units of binary level 4(g-1) in blocks of 16
    prm.ngroups = 0;
    {
        // Geometry: a side range holds about ns/2 elements = R leaves.  The chain levels
        // together climb T = log2(R) - 1 binary levels (so that the top units are R/2..R/4
        // leaves and what is left for the binary walk is 2-4 units), split evenly over as
        // few levels as keep a block at <= 2^16 units.
        int bits[kMaxGroups + 1] = {0, 0, 0, 0, 0};
        int G = 0;
        int T = 0;
        for (int64_t r = (int64_t)h->ns / (2 * kLeaf); r > 1; r >>= 1) ++T;
        T -= 1;
        // delta > 0 with few fitted points: the chains would cost more to build than they save
        const bool sparse = h->delta > 0.0 && (int64_t)h->nfit * 64 < (int64_t)h->n;
        if (T >= 3 && !sparse) {
            // one chain level of up to 16 bits (measured at n = 1e8, ns = 5e6: 16 -> 118.9 ms per
            // smooth, 8 + 8 -> 120.0, 15 -> 125.9, 12 + 4 -> 132.2: what counts is that the top
            // units are R/2..R/4 leaves, and every further level costs two more nodes per range)
            constexpr int kMaxBits = 16;
            G = (T + kMaxBits - 1) / kMaxBits;
            if (G > kMaxGroups) G = kMaxGroups;
            for (int g = 1; g <= G; ++g) bits[g] = T / G + (g <= T % G ? 1 : 0);
            for (int g = 1; g <= G; ++g) if (bits[g] > kMaxBits) bits[g] = kMaxBits;
        }
        if (const char* v = dev_env("CF_TREE_BITS")) {        // development: "8,8" etc.
            G = 0;
            while (*v && G < kMaxGroups) {
                bits[++G] = atoi(v);
                while (*v && *v != ',') ++v;
                if (*v == ',') ++v;
            }
        }
        prm.ulev[1] = 0;
        for (int g = 1; g <= G; ++g) {
            const int ul = prm.ulev[g];
            if (bits[g] < 1 || ul >= L || prm.cnt[ul] <= (1 << bits[g])) break;
            prm.cbits[g] = bits[g];
            prm.ulev[g + 1] = ul + bits[g];
            prm.coff[g] = total;
            total += 2 * (int64_t)prm.cnt[ul];
            prm.ngroups = g;
        }
    }
    prm.total = total;
    // the part of the tree inside each block of 2^cb leaves is built by one launch (with the level-1
    // chain when its blocks are those); CF_TREE_FUSED=0 (development): one launch per level as before
    static const bool fused = !(dev_env("CF_TREE_FUSED") && atoi(dev_env("CF_TREE_FUSED")) == 0);
    // the level-1 chain is built inside the block launch when its blocks are small enough to be
    // the launch's blocks; wider chain blocks get their own launch after the binary levels
    const bool chain_in_block = prm.ngroups >= 1 && prm.cbits[1] <= 9;
    int cb = chain_in_block ? prm.cbits[1] : 8;
    if (cb > L - 1) cb = L - 1;
    prm.qbits = (fused && cb >= 5) ? cb : 0;           // core prefix sums per block of that launch
    const size_t node_doubles = kNodeLen * ((size_t)total * (size_t)h->nseries + kStage);
    const size_t qsum_doubles = prm.qbits ? (size_t)kQLen * prm.cnt[0] * (size_t)h->nseries : 0;
    const size_t qhdr_doubles = prm.qbits ? 2 * (size_t)((prm.cnt[0] >> prm.qbits) + 1) * (size_t)h->nseries : 0;
    // + kStage nodes: the evaluator stages kStage consecutive nodes from any index it needs
    const size_t need = sizeof(double) * (node_doubles + qsum_doubles + qhdr_doubles);
    if (need > h->tree_bytes) {
        cudaFree(h->tree);
        h->tree = nullptr;
        h->tree_bytes = 0;
        CF_CUDA(cudaMalloc(&h->tree, need));
        h->tree_bytes = need;
    }
    if (h->reserve_only) return CF_OK;
    prm.nodes = h->tree;
    prm.qsum = h->tree + node_doubles;
    prm.qhdr = prm.qsum + qsum_doubles;
    prm.span = h->build_span ? h->build_span : h->span;       // what the build is gated to
    TreeParams ptop = prm;                                     // the top levels: everything that has arrived
    if (h->build_span) ptop.span = h->build_span + 2;
    auto mark = [&]() {
        if (!h->timing) return;
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, h->stream);
        h->marks.push_back(e);
    };
    mark();
    if (!h->tree_skip_build) {     // api.cu's chunked last pass: the tree of this pass is already built
        int level = 1;
        int g0 = 1;
        if (fused) {
            const bool chain = chain_in_block;
            // few blocks: 128 threads per block (shortest critical path); many: one warp per block
            static const int bt_env = dev_env("CF_TREE_BLOCK") ? atoi(dev_env("CF_TREE_BLOCK")) : 0;
            const int64_t leaves = (int64_t)prm.cnt[0] * h->nseries;
            const int bt = bt_env ? bt_env : (leaves < (1 << 20) ? 128 : 32);
            const dim3 bg((prm.cnt[0] + (1 << cb) - 1) >> cb, h->nseries);
            int seg = bt == 128 ? 8 : 16;                            // units per chain task
            if (seg > (1 << cb)) seg = 1 << cb;
            if (bt == 128) tree_block_kernel<128><<<bg, 128, 0, h->stream>>>(prm, cb, chain, seg);
            else           tree_block_kernel<32><<<bg, 32, 0, h->stream>>>(prm, cb, chain, seg);
            CF_CUDA(cudaGetLastError());
            h->launches++;
            level = cb + 1;
            if (chain) g0 = 2;
        } else {
            tree_leaf_kernel<<<dim3((prm.cnt[0] + 127) / 128, h->nseries), 128, 0, h->stream>>>(prm);
            CF_CUDA(cudaGetLastError());
            h->launches++;
        }
        for (; level < L && prm.cnt[level] > 256; ++level) {     // one launch per level while it fills more than a CTA
            tree_merge_kernel<<<dim3((prm.cnt[level] + 127) / 128, h->nseries), 128, 0, h->stream>>>(prm, level);
            CF_CUDA(cudaGetLastError());
            h->launches++;
        }
        if (level < L) {
            tree_top_kernel<<<h->nseries, 256, 0, h->stream>>>(ptop, level);
            CF_CUDA(cudaGetLastError());
            h->launches++;
        }
        for (int g = g0; g <= prm.ngroups; ++g) {
            const int32_t ucnt = prm.cnt[prm.ulev[g]];
            const int32_t seg = (1 << prm.cbits[g]) < kChainSeg ? (1 << prm.cbits[g]) : kChainSeg;
            const int32_t nseg = (ucnt + seg - 1) / seg;
            if ((int64_t)nseg * h->nseries >= 65536)      // enough segments to fill the GPU: 1 merge per entry
                tree_chain_kernel<<<dim3((2 * nseg + 127) / 128, h->nseries), 128, 0, h->stream>>>(prm, g);
            else                                          // short critical path instead
                tree_chain_entry_kernel<<<dim3((2 * ucnt + 127) / 128, h->nseries), 128, 0, h->stream>>>(prm, g);
            CF_CUDA(cudaGetLastError());
            h->launches++;
        }
    }
    mark();
    if (h->tree_build_only) return CF_OK;
    const int cntp = prm.end - prm.begin;
    // 7 CTAs of 128 threads per SM (72 registers), tricube nodes staged per warp, core nodes read
    // directly; the others are kept for measurement in development builds (B200, n = 1e8 smooth,
    // round 1: 5 CTAs 143.2 ms, 6: 128.6, 7: 125.7, 7 + L1 prefetch 125.3; round 2, one 16-bit
    // chain level: staged tricube + core nodes 111.7, core nodes direct 108.9)
    static const int variant = dev_env("CF_TREE_EVAL") ? atoi(dev_env("CF_TREE_EVAL")) : 73;
    const dim3 eg((cntp + 127) / 128, h->nseries);
    switch (variant) {
#ifdef CF_DEVELOPMENT
    case 5:  tree_eval_kernel<128, 5, 0><<<eg, 128, 0, h->stream>>>(prm); break;
    case 6:  tree_eval_kernel<128, 6, 0><<<eg, 128, 0, h->stream>>>(prm); break;
    case 70: tree_eval_kernel<128, 7, 0><<<eg, 128, 0, h->stream>>>(prm); break;
    case 71: tree_eval_kernel<128, 7, 1><<<eg, 128, 0, h->stream>>>(prm); break;   // round 1: per-lane node loads
    case 62: tree_eval_kernel<128, 6, 2><<<eg, 128, 0, h->stream>>>(prm); break;
    case 52: tree_eval_kernel<128, 5, 2><<<eg, 128, 0, h->stream>>>(prm); break;

    case 72: tree_eval_kernel<128, 7, 2><<<eg, 128, 0, h->stream>>>(prm); break;   // core nodes staged too
    case 74: tree_eval_kernel<128, 7, 4><<<eg, 128, 0, h->stream>>>(prm); break;   // accounting: phase 2 skipped (wrong results)
    case 75: tree_eval_kernel<128, 7, 5><<<eg, 128, 0, h->stream>>>(prm); break;   // accounting: phase 3 skipped (wrong results)
#endif
    default: tree_eval_kernel<128, 7, 3><<<eg, 128, 0, h->stream>>>(prm); break;   // warp-staged tricube nodes
    }
    CF_CUDA(cudaGetLastError());
    h->launches++;
    h->fit_launches++;
    return CF_OK;
}

}  // namespace cf
