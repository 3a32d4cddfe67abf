// common.cuh -- internal declarations shared by the libcurvefit_cuda sources.
// sm_100a only.  Nothing here is part of the public C ABI (include/curvefit_cuda.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <vector>

#include "curvefit_cuda.h"

namespace cf {

// Development switches (tile variants, tree geometry sweeps) are read from the environment
// only in builds made with -DCF_DEVELOPMENT; a release build has no environment routes.
#ifdef CF_DEVELOPMENT
#include <stdlib.h>
inline const char* dev_env(const char* name) { return getenv(name); }
#else
inline const char* dev_env(const char*) { return nullptr; }
#endif

// ---- error plumbing ---------------------------------------------------------
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define CF_CUDA(call)                                                         \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess)                                               \
            return ::cf::cuda_fail(e__, #call, __FILE__, __LINE__);           \
    } while (0)

#define CF_TRY(call)                                                          \
    do {                                                                      \
        int rc__ = (call);                                                    \
        if (rc__ != CF_OK) return rc__;                                       \
    } while (0)

// ---- select (median of |res|) geometry ---------------------------------------
// Keys are the bit patterns of |res| (non-negative doubles order as uint64; bit 63 is
// always clear, so TOP = 63 informative bits) or, for cf_median, the order-preserving
// image of the signed value (TOP = 64).  They are resolved most-significant digit first:
// round 0 takes the top B0 bits of EVERY element (fused into the residual pass), later
// rounds take W bits each of the CANDIDATES only (robust.cu).  Two ranks are tracked
// together: the reference reads rw(m1) and rw(m2), m1 = 1 + n/2, m2 = n - m1 + 1
// (src/curvefit_regression.f90:699-701).
constexpr int kSelB0Single = 15, kSelWSingle = 12;     // one series: 128 KB shared-memory histogram
constexpr int kSelB0Batched = 11, kSelWBatched = 11;   // nseries > 1: per-series state stays small

struct SelGeom { int top, b0, w, rounds; };            // rounds = number of W-bit rounds after round 0
__host__ __device__ inline SelGeom sel_geom(bool ordered, bool batched)
{
    SelGeom g;
    g.top = ordered ? 64 : 63;
    g.b0 = batched ? kSelB0Batched : kSelB0Single;
    g.w = batched ? kSelWBatched : kSelWSingle;
    g.rounds = (g.top - g.b0 + g.w - 1) / g.w;
    return g;
}
// round k = 1 .. rounds resolves bits [shift, shift + width)
__host__ __device__ inline int sel_width(const SelGeom& g, int k)
{
    return k == 1 ? (g.top - g.b0) - g.w * (g.rounds - 1) : g.w;
}
__host__ __device__ inline int sel_shift(const SelGeom& g, int k) { return g.w * (g.rounds - k); }

struct SelectState {
    unsigned long long prefix[2];  // resolved high bits of the two order statistics
    unsigned long long rank[2];    // remaining 0-based rank inside the prefix bucket
    double cmad, c1, c9;           // :701-703
    unsigned int bucket[2];        // round 0: top-B0-bit bucket of each tracked rank
    unsigned int ncand;            // keys appended to the candidate list (compact_kernel)
    unsigned int ticket;           // CTAs of the running select kernel that have finished
};

#ifdef __CUDACC__
// order-preserving uint64 image of a double (negative values: all bits flipped;
// non-negative: sign bit set) and its inverse
__device__ __forceinline__ unsigned long long ordered_key(double v)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double ordered_value(unsigned long long k)
{
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    return __longlong_as_double((long long)b);
}

// bisquare robustness weight (:704-713), IEEE operations in the reference's order
__device__ __forceinline__ double bisquare_weight(double res, double cmad, double c1, double c9)
{
    const double r = fabs(res);
    if (r <= c1) return 1.0;
    if (r > c9) return 0.0;
    const double t = __ddiv_rn(r, cmad);
    const double u = __dsub_rn(1.0, __dmul_rn(t, t));
    return __dmul_rn(u, u);
}

// local linear fit evaluated at xs from the five moments (:559-584)
__device__ __forceinline__ double fit_value(double S0, double S1, double S2, double T0, double T1,
                                            double yi, double h, double scale, double us,
                                            double range)
{
    if (!(S0 > 0.0)) return yi;                       // ok = .false., :559, :673
    const double inv = 1.0 / S0;
    const double ym = T0 * inv;                       // weighted mean of y
    double out = ym;
    if (h > 0.0) {                                    // :564
        const double um = S1 * inv;
        double cu = fma(-um, um, S2 * inv);           // weighted variance of u
        cu = cu > 0.0 ? cu : 0.0;
        if (sqrt(cu) * scale > 0.001 * range) {       // :573 (c = cu * scale^2)
            const double cov = fma(-um, ym, T1 * inv);
            out = fma(cov / cu, us - um, ym);
        }
    }
    return out;
}
#endif

// ---- peer-memory exchange (peer.cu): the robustness step's two collectives done
// by the kernels themselves over NVLink peer stores instead of NCCL ------------
constexpr int kMaxPeers = 8;
struct PeerComm {                                   // one per rank, mapped by all peers
    unsigned long long flag[2][kMaxPeers];          // [0] histogram all-reduce, [1] barriers: epoch per source rank
    int error;                                      // set when a wait timed out
    int32_t span[2];                                // this rank's element span (windows.cu), published
    int32_t peer_span[kMaxPeers][2];                // every rank's span, local copy made after the barrier
    unsigned int cand_count[kMaxPeers];             // candidates rank r pushed into slot r of this rank's list
    unsigned int slots[2][kMaxPeers][1 << kSelB0Single];  // [epoch parity][source rank] round-0 histograms
};
struct PeerPtrs {                                   // by-value kernel argument
    PeerComm* comm[kMaxPeers];
    double* rw[kMaxPeers];
    double* ys[kMaxPeers];
    double* res[kMaxPeers];
    unsigned long long* cand[kMaxPeers];            // candidate lists (nranks slots of `block` keys)
    unsigned long long timeout_ns;                  // device-side waits give up after this long
    int rank, nranks;
};

// ---- per-point neighbourhood (0-based, inclusive) ----------------------------
struct Windows {
    int32_t* nleft;   // slide result, :663-669
    int32_t* lo9;     // first j >= nleft with |x_j - xs| <= .999 h   (:546)
    int32_t* hi9;     // last  j          with |x_j - xs| <= .999 h
    int32_t* lo1;     // first j in [lo9,hi9] with |x_j - xs| <= .001 h (w = 1 core)
    int32_t* hi1;     // last  such j
    int32_t* leader;  // first index of the tie group of i (:683-689)
};

}  // namespace cf

// ---- the handle ---------------------------------------------------------------
struct cf_lowess {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev[8] = {};

    int32_t n = 0;
    int32_t nseries = 1;   // independent series of length n stored back to back (batched mode)
    bool init = false;

    // shard of the sorted index range fitted by this handle
    int32_t rank = 0, nranks = 1, block = 0, begin = 0, end = 0;

    // device buffers (capacity in elements)
    int64_t cap = 0, cap_rw = 0;
    double *x = nullptr, *y = nullptr, *ys = nullptr, *rw = nullptr, *res = nullptr;
    cf::Windows win = {};   // six int32 arrays carved from ONE block (win.nleft = its start): the
                            // block doubles as the radix sort's 24 n bytes of (key, index) buffers
    void* sort_aux = nullptr;            // digit table of the sort, kept between calls
    size_t sort_aux_cap = 0;
    double win_f = -1.0;  // span fraction the stored windows were built for
    int32_t ns = 0;

    // select (robust.cu): per series the round-0 histogram, the two W-bit histograms, the state
    unsigned int* hist0 = nullptr;       // nseries << B0 counters, zero between uses
    unsigned int* hist = nullptr;        // nseries x 2 << W counters, zero between uses
    cf::SelectState* sel = nullptr;      // nseries states
    int32_t sel_cap = 0;
    unsigned long long* cand = nullptr;  // candidate keys: nseries x cand_stride
    int64_t cap_cand = 0, cand_stride = 0;
    bool sel_attr = false;               // dynamic shared-memory opt-in done on this device
    int32_t* span = nullptr;             // nseries x {min lo9, max hi9} over the shard (windows.cu)

    // fit kernel: partial-moment scratch for the K-way window split, tuning knobs
    double* part = nullptr;
    size_t part_bytes = 0;
    int sms = 0;
    int fit_variant = 0;       // env CF_FIT_VARIANT (development only)
    int ksplit_override = 0;   // env CF_FIT_KSPLIT  (development only)
    // cf_lowess_fit_sample: launch only tiles tile_first + k * tile_stride, k < tile_count (0 = all)
    int32_t tile_first = 0, tile_stride = 1, tile_count = 0, tile_ksplit = 1;

    // moment-tree evaluator (tree.cu): node storage, rebuilt every pass
    double* tree = nullptr;
    size_t tree_bytes = 0;
    int method = 0;            // CF_METHOD_DIRECT / CF_METHOD_TREE / CF_METHOD_AUTO

    // delta > 0 (delta.cu): lowest() runs only at the fitted points fidx[0..nfit), the
    // rest is tie-copied / interpolated (src/curvefit_regression.f90:674-690)
    double delta = 0.0;
    double delta_built = -1.0;           // delta the lists below were built for
    bool x_partial = false;              // x holds -inf / +inf sentinels outside an uploaded span (upload_span):
                                         // the fitted-point list of delta > 0 cannot be derived from it
    int32_t nfit = 0;
    int32_t slot_begin = 0, slot_end = 0; // slots of fidx this shard fits: its own fitted points and the one
                                          // fitted point before / after its block (the interpolation needs both)
    int32_t* fidx = nullptr;             // fitted points, ascending
    int32_t* fcend = nullptr;            // end of the tie chain of each fitted point (`last`, :681-689)
    int32_t* frank = nullptr;            // n + 1: number of fitted points < p
    int64_t cap_delta = 0;

    // peer-memory exchange (peer.cu); comm is this rank's own block
    cf::PeerComm* comm = nullptr;
    cf::PeerPtrs peer = {};
    bool peer_attached = false;
    bool peer_ipc = false;               // peers mapped through CUDA IPC (else: same-process peer access)
    unsigned long long peer_timeout_ns = 60000000000ull;
    unsigned long long epoch_hist = 0, epoch_bar = 0;

    // cf_lowess_smooth (host buffers): second stream for copies that overlap kernels -- y's upload
    // runs beside the window search, the download of the final robustness weights beside the
    // last fit pass (they are final once the last robustness update is done, :704-713)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy[2] = {};
    double* early_rw = nullptr;           // host destination of that download (null: none), full-length base
    double* early_rw_missed = nullptr;    // a requested early download that the last run could NOT issue: whoever
                                          // fetches the results copies the weights there instead
    double* early_ys = nullptr;           // cf_lowess_smooth: host destinations of ys / res, downloaded chunk by
    double* early_res = nullptr;          //   chunk while the last pass of the tree evaluator is still running
    bool reserve_only = false;            // launch_fit: size and allocate the evaluator's scratch (partial moments /
                                          // tree nodes), launch nothing -- multi.cu does this before its host barrier
    bool tree_skip_build = false;         // tree.cu: evaluate only (the pass's tree is built)
    bool tree_build_only = false;         // tree.cu: build only
    // pipelined first pass of cf_lowess_smooth (api.cu): the tree is built chunk by chunk as the
    // elements arrive; build_span[0..1] gates the block / merge / chain kernels to the new chunk,
    // build_span[2..3] the top levels to everything that has arrived (device ints; null: h->span)
    int32_t* build_span = nullptr;        // active only while a chunk is being built
    int32_t* build_span_buf = nullptr;    // its allocation
    std::vector<cudaEvent_t> chunk_events;
    bool first_fit_done = false;          // run_passes: the first fit pass was already issued
    bool last_in_pieces = false;          // run_passes evaluated the last pass in pieces and sent ys / res home
    bool y_pending = false;               // y's upload is in flight on copy_stream (ev_copy[1])

    // counters for cf_stats
    int launches = 0, fit_launches = 0;
    bool timing = false;                  // cf_lowess_run with stats: stage events are recorded
    std::vector<cudaEvent_t> marks;       // pairs (start, end) around each tree rebuild
};

namespace cf {

// kernels' host launchers (each returns CF_OK or an error code; async on h->stream)
int launch_windows(cf_lowess* h, int32_t ns);
int launch_windows_range(cf_lowess* h, int32_t ns, int32_t begin, int32_t end, bool init_span);   // a sub-range of points
int launch_fit(cf_lowess* h, bool use_rw);        // dispatches on h->method
int launch_fit_tree(cf_lowess* h, bool use_rw);   // tree.cu
bool method_is_tree(const cf_lowess* h);
int launch_residuals(cf_lowess* h);           // tie copy + res = y - ys on the shard
int delta_prepare(cf_lowess* h);              // delta.cu: (re)build the fitted-point list
int launch_delta_fill(cf_lowess* h);          // delta.cu: tie copy + interpolation + res
void delta_free(cf_lowess* h);
// robust.cu: the select of the two middle order statistics of |res| and the weights
int stateless_method();                                    // api.cu: evaluator of this thread's stateless call
int select_prepare(cf_lowess* h);                          // (re)allocate the candidate list
int select_rounds(const cf_lowess* h);                     // W-bit rounds after round 0
int launch_select(cf_lowess* h, bool fused_residuals);     // passes A-C; single GPU or peer-attached
int launch_hist0(cf_lowess* h, bool fused, bool defer);    // pass A (defer: counters stay in hist0)
int launch_pick0(cf_lowess* h);                            // round 0 scan on its own
int launch_compact(cf_lowess* h);                          // pass B (+ peer exchange of the candidates)
int launch_cand_round(cf_lowess* h, int k, bool defer);    // pass C, round k
int launch_cand_pick(cf_lowess* h, int k);
int launch_bisquare(cf_lowess* h);
size_t select_scratch_bytes(int32_t n);
int select_median(const double* d_x, int32_t n, double* d_out, void* scratch, cudaStream_t s, int* launches);
int launch_peer_push(cf_lowess* h);           // peer.cu: own candidates -> every rank's list, then a barrier
int launch_peer_barrier(cf_lowess* h);        // peer.cu: all ranks reach this point
int launch_peer_gather(cf_lowess* h, int what);  // peer.cu: own ys/res/rw block -> every rank
int launch_peer_spans(cf_lowess* h);          // peer.cu: publish the shard's element span, learn the peers'
void peer_detach(cf_lowess* h);               // peer.cu
int peer_prepare(cf_lowess* h);               // (re)allocate and clear what the peers map
int peer_check(cf_lowess* h);                 // CF_DEVICE_ERROR when a device-side wait timed out
int peer_clear_error(cf_lowess* h);

// device radix sort of (x, index) ascending, y co-permuted; d_perm may be null
int sort_xy_device(const double* d_x_in, const double* d_y_in, int32_t n, double* d_x_out,
                   double* d_y_out, int32_t* d_perm_out, cudaStream_t s, int* launches,
                   void* pairs = nullptr, void* aux = nullptr);
size_t sort_aux_bytes(int32_t n);
int sort_xy_inplace(double* x_io, double* y_io, int32_t n, double* x_out, double* y_out,
                    void* scratch16n, void* aux, cudaStream_t s, int* launches);

int fp64_peak(int device, double ms_target, double* tflops, double* mhz);

// api.cu, used by multi.cu (cf_lowess_smooth_multi): allocate for n points without uploading;
// upload elements [lo, hi] only (x outside becomes -inf / +inf sentinels)
int ensure_copy_stream(cf_lowess* h);
int layout_series(cf_lowess* h, int32_t n);
int upload_span(cf_lowess* h, const double* x, const double* y, int32_t lo, int32_t hi);
void shard_span_host(const double* x, int32_t n, int32_t ns, int32_t begin, int32_t end, int32_t* lo_out,
                     int32_t* hi_out);
const char* last_error_text();

}  // namespace cf
