// api.cu -- the C ABI of libcurvefit_cuda (include/curvefit_cuda.h): handle
// lifecycle, the lowess driver loop (src/curvefit_regression.f90:657-714 as a
// chain of device stages), and the stateless host-buffer entry points the
// Fortran shim binds.  No CPU fallback: every entry point that does work needs
// a CUDA device and reports CF_DEVICE_ERROR otherwise.
#include <math.h>
#include <math_constants.h>
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <new>
#include <vector>

#include <nvtx3/nvToolsExt.h>      // header-only: ranges cost nothing unless a profiler is attached

#include "common.cuh"

namespace cf {

// NVTX range over a host-side stage (launch sequence) of the lowess driver loop
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char* last_error_text() { return g_err; }

int cuda_fail(cudaError_t e, const char* what, const char* file, int line)
{
    set_error("CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
    return e == cudaErrorMemoryAllocation ? CF_OUT_OF_MEMORY_ERROR : CF_DEVICE_ERROR;
}

constexpr int kTile = 512;  // fit.cu's T = NT*P; shard blocks are multiples of it

static void compute_shard(cf_lowess* h)
{
    const int64_t n = h->n;
    int64_t per = (n + h->nranks - 1) / h->nranks;
    per = (per + kTile - 1) / kTile * kTile;
    if (per < kTile) per = kTile;
    h->block = (int32_t)per;
    const int64_t b = (int64_t)h->rank * per;
    h->begin = (int32_t)(b < n ? b : n);
    h->end = (int32_t)(b + per < n ? b + per : n);
}

// Exported (IPC-mapped) buffers are about to be freed: the mapping of the peers'
// buffers is dropped here; the peers must export / attach again as well.
static void free_buffers(cf_lowess* h)
{
    peer_detach(h);
    cudaFree(h->x); cudaFree(h->y); cudaFree(h->ys); cudaFree(h->rw); cudaFree(h->res);
    cudaFree(h->win.nleft);      // one block holds all six window arrays
    h->x = h->y = h->ys = h->rw = h->res = nullptr;
    h->win = Windows{};
    h->cap = h->cap_rw = 0;
}

static int reserve(cf_lowess* h, int32_t n, int32_t nseries = 1)
{
    const int64_t need = (n > 0 ? (int64_t)n : 1) * (nseries > 0 ? nseries : 1);
    if (nseries > h->sel_cap) {
        cudaFree(h->hist0); cudaFree(h->hist); cudaFree(h->sel); cudaFree(h->span);
        h->hist0 = nullptr; h->hist = nullptr; h->sel = nullptr; h->span = nullptr; h->sel_cap = 0;
        // sized for either geometry (one series: B0 = 15, W = 12; batched: 11, 11)
        size_t b0 = (sizeof(unsigned int) << kSelB0Batched) * (size_t)nseries;
        size_t bw = 2 * (sizeof(unsigned int) << kSelWBatched) * (size_t)nseries;
        if (b0 < (sizeof(unsigned int) << kSelB0Single)) b0 = sizeof(unsigned int) << kSelB0Single;
        if (bw < 2 * (sizeof(unsigned int) << kSelWSingle)) bw = 2 * (sizeof(unsigned int) << kSelWSingle);
        CF_CUDA(cudaMalloc(&h->span, sizeof(int32_t) * 2 * (size_t)nseries));
        CF_CUDA(cudaMalloc(&h->hist0, b0));
        CF_CUDA(cudaMalloc(&h->hist, bw));
        CF_CUDA(cudaMalloc(&h->sel, sizeof(SelectState) * (size_t)nseries));
        h->sel_cap = nseries;
    }
    {   // the select kernels leave their counters zeroed; start from zero whatever happened before
        const SelGeom g = sel_geom(false, nseries > 1);
        CF_CUDA(cudaMemsetAsync(h->hist0, 0, (sizeof(unsigned int) << g.b0) * (size_t)nseries, h->stream));
        CF_CUDA(cudaMemsetAsync(h->hist, 0, 2 * (sizeof(unsigned int) << g.w) * (size_t)nseries, h->stream));
        CF_CUDA(cudaMemsetAsync(h->sel, 0, sizeof(SelectState) * (size_t)nseries, h->stream));
    }
    if (need > h->cap) {
        if (h->peer_attached) {
            set_error("the series grew while peers are attached: detach on all ranks first");
            return CF_INVALID_INPUT_ERROR;
        }
        free_buffers(h);
        const size_t bd = sizeof(double) * (size_t)need, bi = sizeof(int32_t) * (size_t)need;
        CF_CUDA(cudaMalloc(&h->x, bd));
        CF_CUDA(cudaMalloc(&h->y, bd));
        CF_CUDA(cudaMalloc(&h->ys, bd));
        CF_CUDA(cudaMalloc(&h->res, bd));
        CF_CUDA(cudaMalloc(&h->win.nleft, 6 * bi));
        h->win.lo9 = h->win.nleft + need;
        h->win.hi9 = h->win.lo9 + need;
        h->win.lo1 = h->win.hi9 + need;
        h->win.hi1 = h->win.lo1 + need;
        h->win.leader = h->win.hi1 + need;
        h->cap = need;
    }
    return CF_OK;
}

// rw is padded to nranks * block so that an all-gather of equal blocks fits
static int reserve_rw(cf_lowess* h)
{
    int64_t need = (int64_t)h->nranks * h->block;
    if (need < h->n) need = h->n;
    if (h->nseries > 1) need = (int64_t)h->n * h->nseries;   // batched: no sharding inside a series
    if (need < 1) need = 1;
    if (need > h->cap_rw) {
        if (h->peer_attached) {
            set_error("the series grew while peers are attached: detach on all ranks first");
            return CF_INVALID_INPUT_ERROR;
        }
        cudaFree(h->rw);
        h->rw = nullptr;
        h->cap_rw = 0;
        CF_CUDA(cudaMalloc(&h->rw, sizeof(double) * (size_t)need));
        CF_CUDA(cudaMemsetAsync(h->rw, 0, sizeof(double) * (size_t)need, h->stream));
        h->cap_rw = need;
    }
    return CF_OK;
}

static int use_device(const cf_lowess* h)
{
    CF_CUDA(cudaSetDevice(h->device));
    return CF_OK;
}

// residuals (:695) fused with round 0 of the select, the select (:697-703), the weights (:704-713)
static int residuals_and_robust_update(cf_lowess* h)
{
    const bool fused = !(h->delta > 0.0);
    if (!fused) CF_TRY(launch_residuals(h));       // delta > 0: interpolation + tie copy + residuals
    CF_TRY(launch_select(h, fused));
    return launch_bisquare(h);
}

// upload_span leaves y's copy in flight on the copy stream: whatever touches y next on the
// handle's stream waits for it first
static int wait_pending_y(cf_lowess* h)
{
    if (h->y_pending) {
        CF_CUDA(cudaStreamWaitEvent(h->stream, h->ev_copy[1], 0));
        h->y_pending = false;
    }
    return CF_OK;
}

static int begin_stage(cf_lowess* h, double f)
{
    if (!h->init) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    if (!(f == f)) { set_error("f is NaN"); return CF_INVALID_INPUT_ERROR; }
    const int32_t ns = cf_lowess_span(h->n, f);
    if (h->n >= 2 && h->delta > 0.0 && h->delta_built != h->delta) {
        CF_TRY(delta_prepare(h));          // fitted-point list (src/curvefit_regression.f90:681-690)
        h->win_f = -1.0;                   // windows are per fitted point
    }
    if (h->n >= 2 && (h->win_f != f || h->ns != ns)) {
        NvtxRange r("windows");
        h->ns = ns;
        CF_TRY(launch_windows(h, ns));
        h->win_f = f;
    }
    h->ns = ns;
    return wait_pending_y(h);              // upload_span: y arrives on the copy stream
}

// evaluator of handles created from now on and of the stateless calls (cf_set_default_method)
static std::atomic<int> g_default_method{CF_METHOD_DIRECT};
// cf_lowess_smooth_ex: evaluator of the stateless call this THREAD is inside (-1: the process default)
static thread_local int t_call_method = -1;
int stateless_method() { return t_call_method >= 0 ? t_call_method : g_default_method.load(); }

}  // namespace cf

using namespace cf;

extern "C" {

int32_t cf_version(void) { return 100; }

int32_t cf_last_error(char* buf, int32_t len)
{
    const int32_t need = (int32_t)strlen(g_err);
    if (buf && len > 0) {
        strncpy(buf, g_err, (size_t)len - 1);
        buf[len - 1] = '\0';
    }
    return need;
}

int32_t cf_lowess_span(int32_t n, double f)
{
    // ns = max(min(int(f * real(n)), n), 2), src/curvefit_regression.f90:647
    const double fn = f * (double)n;
    int32_t k;
    if (!(fn < 2147483647.0)) k = n;
    else if (fn < -2147483647.0) k = 2;
    else k = (int32_t)fn;
    if (k > n) k = n;
    if (k < 2) k = 2;
    return k;
}

int32_t cf_lowess_select_rounds(void) { return 1 + sel_geom(false, false).rounds; }
int32_t cf_lowess_hist0_len(void) { return 1 << kSelB0Single; }
int32_t cf_lowess_hist_len(void) { return 2 << kSelWSingle; }

int32_t cf_lowess_create(cf_lowess** out, int32_t device, void* stream)
{
    if (!out) { set_error("null handle pointer"); return CF_INVALID_INPUT_ERROR; }
    *out = nullptr;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available (%s); libcurvefit_cuda has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return CF_DEVICE_ERROR;
    }
    if (device < 0) CF_CUDA(cudaGetDevice(&device));
    if (device >= ndev) { set_error("device %d out of range (%d)", device, ndev); return CF_INVALID_INPUT_ERROR; }
    CF_CUDA(cudaSetDevice(device));
    cf_lowess* h = new (std::nothrow) cf_lowess();
    if (!h) { set_error("Insufficient memory available."); return CF_OUT_OF_MEMORY_ERROR; }
    h->device = device;
    if (const char* v = dev_env("CF_FIT_VARIANT")) h->fit_variant = atoi(v);
    if (const char* v = dev_env("CF_FIT_KSPLIT")) h->ksplit_override = atoi(v);
    h->method = g_default_method.load();
    if (stream) {
        h->stream = (cudaStream_t)stream;
    } else {
        e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) { delete h; return cuda_fail(e, "cudaStreamCreate", __FILE__, __LINE__); }
        h->own_stream = true;
    }
    for (auto& ev : h->ev) {
        e = cudaEventCreate(&ev);
        if (e != cudaSuccess) { cf_lowess_destroy(h); return cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__); }
    }
    {
        const int rc = reserve(h, 1, 1);   // select state + histograms for one series
        if (rc != CF_OK) { cf_lowess_destroy(h); return rc; }
    }
    *out = h;
    return CF_OK;
}

void cf_lowess_destroy(cf_lowess* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    free_buffers(h);
    cudaFree(h->comm);
    cudaFree(h->hist0);
    cudaFree(h->hist);
    cudaFree(h->cand);
    cudaFree(h->sel);
    cudaFree(h->span);
    cudaFree(h->sort_aux);
    cudaFree(h->part);
    cudaFree(h->tree);
    delta_free(h);
    for (auto& ev : h->ev) if (ev) cudaEventDestroy(ev);
    for (auto& ev : h->ev_copy) if (ev) cudaEventDestroy(ev);
    for (cudaEvent_t ev : h->chunk_events) cudaEventDestroy(ev);
    cudaFree(h->build_span_buf);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

static int32_t init_common(cf_lowess* h, const void* x, const void* y, int32_t n, int32_t sort,
                           cudaMemcpyKind kind)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (n < 0 || (n > 0 && (!x || !y))) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    CF_TRY(use_device(h));
    CF_TRY(wait_pending_y(h));                         // an earlier span upload must not land after this one
    h->init = false;                                   // :758
    h->win_f = -1.0;
    h->delta_built = -1.0;
    h->x_partial = false;
    h->nseries = 1;
    CF_TRY(reserve(h, n));
    h->n = n;
    compute_shard(h);
    CF_TRY(reserve_rw(h));
    const size_t bytes = sizeof(double) * (size_t)n;
    if (n > 0) {
        if (sort) {
            // stage the unsorted input in ys/res, sort into x/y  (:793-798)
            CF_CUDA(cudaMemcpyAsync(h->ys, x, bytes, kind, h->stream));
            CF_CUDA(cudaMemcpyAsync(h->res, y, bytes, kind, h->stream));
            // scratch: the (not yet computed) window arrays hold the key / index buffers
            const size_t aux = sort_aux_bytes(n);
            if (aux > h->sort_aux_cap) {
                cudaFree(h->sort_aux);
                h->sort_aux = nullptr;
                h->sort_aux_cap = 0;
                CF_CUDA(cudaMalloc(&h->sort_aux, aux));
                h->sort_aux_cap = aux;
            }
            CF_TRY(sort_xy_inplace(h->ys, h->res, n, h->x, h->y, h->win.nleft, h->sort_aux, h->stream,
                                   &h->launches));
        } else {
            CF_CUDA(cudaMemcpyAsync(h->x, x, bytes, kind, h->stream));   // :800-803
            CF_CUDA(cudaMemcpyAsync(h->y, y, bytes, kind, h->stream));
        }
        CF_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->init = true;                                    // :810
    return CF_OK;
}

int32_t cf_lowess_init(cf_lowess* h, const double* x, const double* y, int32_t n, int32_t sort)
{
    return init_common(h, x, y, n, sort, cudaMemcpyHostToDevice);
}

}  // extern "C"

namespace cf {

// second stream + events of the overlapped copies, created on first use
int ensure_copy_stream(cf_lowess* h)
{
    if (h->copy_stream) return CF_OK;
    CF_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (auto& ev : h->ev_copy) CF_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    return CF_OK;
}

// ls_init's allocation part (:758-789) for a series of n points, nothing uploaded yet
// (multi.cu: a rank of cf_lowess_smooth_multi receives only the elements its windows reach).
int layout_series(cf_lowess* h, int32_t n)
{
    if (!h || n < 1) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    CF_TRY(use_device(h));
    h->init = false;
    h->win_f = -1.0;
    h->delta_built = -1.0;
    h->x_partial = false;
    h->nseries = 1;
    CF_TRY(reserve(h, n));
    h->n = n;
    compute_shard(h);
    CF_TRY(reserve_rw(h));
    h->init = true;
    return CF_OK;
}

// x outside [lo, hi] <- -inf below, +inf above: the series stays ascending, so every search of
// windows.cu that probes beyond the uploaded elements still meets a monotone predicate and ends
// on the same index as on the whole series (the answers lie inside [lo, hi] by construction).
// x(1) and x(n) keep their values: lowest() needs range = x(n) - x(1) (:535), and a probe that
// lands on them sees what it would see on the whole series.
__global__ void __launch_bounds__(256)
sentinel_kernel(double* __restrict__ x, int32_t n, int32_t lo, int32_t hi)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t below = lo > 1 ? lo - 1 : 0;                         // indices 1 .. lo - 1
    const int64_t above = (int64_t)n - 2 - hi > 0 ? (int64_t)n - 2 - hi : 0;   // hi + 1 .. n - 2
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < below + above; k += stride) {
        if (k < below) x[1 + k] = -CUDART_INF;
        else x[(int64_t)hi + 1 + (k - below)] = CUDART_INF;
    }
}

// Elements [lo, hi] of the host series -> the same positions of the handle's x, y (asynchronous
// on the handle's stream); the rest of x becomes sentinels (first and last element excepted),
// the rest of y is left as it is (nothing reads it: fits, tree nodes and residuals touch
// [min lo9, max hi9] and the shard only).
int upload_span(cf_lowess* h, const double* x, const double* y, int32_t lo, int32_t hi)
{
    if (!h || !h->init || lo < 0 || hi >= h->n || lo > hi) { set_error("bad span [%d, %d]", (int)lo, (int)hi); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    h->win_f = -1.0;                                   // new data: new neighbourhoods
    h->delta_built = -1.0;
    h->x_partial = lo > 1 || hi < h->n - 2;            // sentinels are written below exactly then
    const size_t bytes = sizeof(double) * (size_t)(hi - lo + 1);
    CF_TRY(ensure_copy_stream(h));
    CF_CUDA(cudaMemcpyAsync(h->x + lo, x + lo, bytes, cudaMemcpyHostToDevice, h->stream));
    // y follows x over the bus on the second stream while the window search (x only) runs;
    // the first fit pass waits for it (begin_stage)
    CF_CUDA(cudaEventRecord(h->ev_copy[0], h->stream));
    CF_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0));
    CF_CUDA(cudaMemcpyAsync(h->y + lo, y + lo, bytes, cudaMemcpyHostToDevice, h->copy_stream));
    CF_CUDA(cudaEventRecord(h->ev_copy[1], h->copy_stream));
    h->y_pending = true;
    if (lo > 0) CF_CUDA(cudaMemcpyAsync(h->x, x, sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (hi < h->n - 1)
        CF_CUDA(cudaMemcpyAsync(h->x + (h->n - 1), x + (h->n - 1), sizeof(double), cudaMemcpyHostToDevice, h->stream));
    if (lo > 1 || hi < h->n - 2) {
        sentinel_kernel<<<296, 256, 0, h->stream>>>(h->x, h->n, lo, hi);
        CF_CUDA(cudaGetLastError());
        h->launches++;
    }
    return CF_OK;
}

// nleft of point i (0-based) as windows.cu finds it: the first L in [0, n - ns] with
// x(i) - x(L) <= x(L + ns) - x(i)  (src/curvefit_regression.f90:664-666; monotone in L)
static int64_t host_nleft(const double* x, int64_t n, int64_t ns, int64_t i)
{
    const double xs = x[i];
    int64_t lo = 0, hi = n - ns;                       // the predicate holds at n - ns by construction
    while (lo < hi) {
        const int64_t mid = lo + ((hi - lo) >> 1);
        const double d1 = xs - x[mid], d2 = x[mid + ns] - xs;
        if (d1 <= d2) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// The elements the windows of output points [begin, end) reach, computed on the host from the
// sorted abscissas: nleft is monotone in i, so they start at nleft(begin) and end at
// nright(end - 1) = nleft(end - 1) + ns - 1, plus the one element beyond that the searches of
// windows.cu test (x(L + ns), the first excluded point), widened over ties -- a window with h = 0
// takes every point equal to xs (:546).  On a grid that is block + ns / 2 either side: the
// (ns-1)-point halo is the worst case (blocks at the ends of the series).
void shard_span_host(const double* x, int32_t n, int32_t ns, int32_t begin, int32_t end, int32_t* lo_out,
                     int32_t* hi_out)
{
    if (end > n) end = n;
    if (begin >= end) begin = end > 0 ? end - 1 : 0;
    int64_t lo = host_nleft(x, n, ns, begin);
    int64_t hi = host_nleft(x, n, ns, end > 0 ? end - 1 : 0) + ns;       // nright + 1
    if (hi > n - 1) hi = n - 1;
    if (lo > begin) lo = begin;
    if (hi < end - 1) hi = end - 1;
    lo = std::lower_bound(x, x + lo, x[lo]) - x;
    hi = (std::upper_bound(x + hi, x + n, x[hi]) - x) - 1;
    *lo_out = (int32_t)lo;
    *hi_out = (int32_t)hi;
}

}  // namespace cf

extern "C" {

int32_t cf_lowess_shard_span(const cf_lowess* h, const double* x_sorted, int32_t n, double f,
                             int32_t* first, int32_t* last)
{
    if (!h || !x_sorted || n < 1 || !first || !last) { set_error("null argument"); return CF_INVALID_INPUT_ERROR; }
    int64_t per = ((int64_t)n + h->nranks - 1) / h->nranks;
    per = (per + kTile - 1) / kTile * kTile;
    if (per < kTile) per = kTile;
    const int64_t b = (int64_t)h->rank * per;
    const int32_t begin = (int32_t)(b < n ? b : n), end = (int32_t)(b + per < n ? b + per : n);
    int32_t lo = 0, hi = 0;
    shard_span_host(x_sorted, n, cf_lowess_span(n, f), begin < n ? begin : n - 1, end, &lo, &hi);
    *first = lo + 1;
    *last = hi + 1;
    return CF_OK;
}

int32_t cf_lowess_block_span(const double* x_sorted, int32_t n, double f, int32_t begin, int32_t end,
                             int32_t* first, int32_t* last)
{
    if (!x_sorted || n < 1 || !first || !last || begin < 1 || end < begin || end > n || !(f == f)) {
        set_error("cf_lowess_block_span: bad argument");
        return CF_INVALID_INPUT_ERROR;
    }
    int32_t lo = 0, hi = 0;
    shard_span_host(x_sorted, n, cf_lowess_span(n, f), begin - 1, end, &lo, &hi);
    *first = lo + 1;
    *last = hi + 1;
    return CF_OK;
}

int32_t cf_lowess_init_span(cf_lowess* h, const double* x_sorted, const double* y, int32_t n,
                            int32_t first, int32_t last)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (n < 1 || !x_sorted || !y) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (first < 1 || last > n || first > last) { set_error("bad span [%d, %d] of %d", (int)first, (int)last, (int)n); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(layout_series(h, n));
    return upload_span(h, x_sorted, y, first - 1, last - 1);
}

int32_t cf_lowess_set_early_weights(cf_lowess* h, double* rw_host)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    if (rw_host) CF_TRY(ensure_copy_stream(h));
    h->early_rw = rw_host;
    return CF_OK;
}

int32_t cf_lowess_fetch_block(cf_lowess* h, double* ys, double* rw, double* res)
{
    if (!h || !h->init || h->nseries != 1) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    CF_TRY(use_device(h));
    if (!rw && h->early_rw_missed) rw = h->early_rw_missed;   // cf_lowess_set_early_weights, not issued by the run
    h->early_rw_missed = nullptr;
    if (h->end > h->begin) {
        const size_t b = sizeof(double) * (size_t)(h->end - h->begin), o = (size_t)h->begin;
        if (ys) CF_CUDA(cudaMemcpyAsync(ys + o, h->ys + o, b, cudaMemcpyDeviceToHost, h->stream));
        if (rw) CF_CUDA(cudaMemcpyAsync(rw + o, h->rw + o, b, cudaMemcpyDeviceToHost, h->stream));
        if (res) CF_CUDA(cudaMemcpyAsync(res + o, h->res + o, b, cudaMemcpyDeviceToHost, h->stream));
    }
    CF_CUDA(cudaStreamSynchronize(h->stream));
    if (h->copy_stream) CF_CUDA(cudaStreamSynchronize(h->copy_stream));   // an early weights download
    return CF_OK;
}

int32_t cf_lowess_init_device(cf_lowess* h, const void* d_x, const void* d_y, int32_t n, int32_t sort)
{
    return init_common(h, d_x, d_y, n, sort, cudaMemcpyDeviceToDevice);
}

int32_t cf_lowess_init_batched(cf_lowess* h, const double* x, const double* y, int32_t nseries,
                               int32_t n)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (nseries < 1 || nseries > 32768 || n < 0 || (n > 0 && (!x || !y))) {
        set_error("Input array sizes must match.");
        return CF_ARRAY_SIZE_ERROR;
    }
    CF_TRY(use_device(h));
    CF_TRY(wait_pending_y(h));
    h->init = false;
    h->win_f = -1.0;
    h->rank = 0; h->nranks = 1;
    h->x_partial = false;
    CF_TRY(reserve(h, n, nseries));
    h->n = n;
    h->nseries = nseries;
    compute_shard(h);
    CF_TRY(reserve_rw(h));
    const size_t bytes = sizeof(double) * (size_t)n * (size_t)nseries;
    if (n > 0) {
        CF_CUDA(cudaMemcpyAsync(h->x, x, bytes, cudaMemcpyHostToDevice, h->stream));
        CF_CUDA(cudaMemcpyAsync(h->y, y, bytes, cudaMemcpyHostToDevice, h->stream));
        CF_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->init = true;
    return CF_OK;
}

int32_t cf_lowess_set_shard(cf_lowess* h, int32_t rank, int32_t nranks)
{
    if (h && h->nseries > 1 && nranks > 1) { set_error("batched handles are not sharded"); return CF_INVALID_INPUT_ERROR; }
    if (!h || nranks < 1 || rank < 0 || rank >= nranks) { set_error("bad shard %d of %d", rank, nranks); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    if (rank != h->rank || nranks != h->nranks) {
        h->win_f = -1.0;           // windows cover the shard only
        h->delta_built = -1.0;     // and so do the slots of the fitted-point list
        peer_detach(h);            // the peers' block layout changes
    }
    h->rank = rank;
    h->nranks = nranks;
    compute_shard(h);
    return reserve_rw(h);
}

int32_t cf_lowess_shard(const cf_lowess* h, int32_t* begin, int32_t* end, int32_t* block)
{
    if (!h) return CF_INVALID_INPUT_ERROR;
    if (begin) *begin = h->begin;
    if (end) *end = h->end;
    if (block) *block = h->block;
    return CF_OK;
}

int32_t cf_lowess_set_method(cf_lowess* h, int32_t method)
{
    if (!h || method < CF_METHOD_DIRECT || method > CF_METHOD_AUTO) {
        set_error("bad method %d", (int)method);
        return CF_INVALID_INPUT_ERROR;
    }
    h->method = method;
    return CF_OK;
}

int32_t cf_lowess_method(const cf_lowess* h) { return h ? h->method : -1; }

int32_t cf_set_default_method(int32_t method)
{
    if (method < CF_METHOD_DIRECT || method > CF_METHOD_AUTO) {
        set_error("bad method %d", (int)method);
        return CF_INVALID_INPUT_ERROR;
    }
    g_default_method.store(method);
    return CF_OK;
}

int32_t cf_default_method(void) { return g_default_method.load(); }

int32_t cf_lowess_set_delta(cf_lowess* h, double delta)
{
    if (!h || !(delta >= 0.0) || !(delta <= 1.7976931348623157e308)) {
        set_error("delta must be finite and >= 0");
        return CF_INVALID_INPUT_ERROR;
    }
    if (delta > 0.0 && h->nseries != 1) {
        set_error("delta > 0 is supported for a single series (not for batched handles)");
        return CF_INVALID_INPUT_ERROR;
    }
    if (delta > 0.0 && h->x_partial) {
        set_error("delta > 0 needs the whole of x on the handle (cf_lowess_init), not a span upload "
                  "(cf_lowess_init_span): the fitted-point list is derived from every abscissa");
        return CF_INVALID_INPUT_ERROR;
    }
    if (delta != h->delta) h->win_f = -1.0;
    h->delta = delta;
    return CF_OK;
}

double cf_lowess_delta(const cf_lowess* h) { return h ? h->delta : 0.0; }

int32_t cf_lowess_fitted(cf_lowess* h, int32_t* count, int32_t* idx, int32_t cap)
{
    if (!h || !h->init) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    if (!(h->delta > 0.0) || h->delta_built != h->delta) {
        set_error("no fitted-point list: delta = 0 or no pass has run with this delta");
        return CF_INVALID_INPUT_ERROR;
    }
    CF_TRY(use_device(h));
    if (count) *count = h->nfit;
    if (idx && cap > 0) {
        const int32_t m = h->nfit < cap ? h->nfit : cap;
        CF_CUDA(cudaMemcpyAsync(idx, h->fidx, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, h->stream));
        CF_CUDA(cudaStreamSynchronize(h->stream));
        for (int32_t k = 0; k < m; ++k) idx[k] += 1;    // 1-based like every index of this ABI
    }
    return CF_OK;
}

int32_t cf_lowess_reset_windows(cf_lowess* h)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    h->win_f = -1.0;
    return CF_OK;
}

int32_t cf_lowess_count(const cf_lowess* h) { return (h && h->init) ? h->n : 0; }

int64_t cf_lowess_launch_count(const cf_lowess* h) { return h ? (int64_t)h->launches : 0; }

}  // extern "C"

// ---------------------------------------------------------------------------
// the driver loop and the rest of the ABI
// ---------------------------------------------------------------------------
extern "C" {

// The last pass of the tree evaluator in `kLastChunks` pieces of output points: the tree is built
// once, then per piece evaluation + residuals, and the piece's ys / res leave for the host on the
// copy stream while the next piece is evaluated (cf_lowess_smooth with host buffers: at n = 1e8
// the 1.6 GB download takes longer than the pass).  A piece is a shard of the shard: per-point
// sums are those of the undivided pass, bit for bit.
constexpr int kLastChunks = 8;
static int last_pass_chunked(cf_lowess* h, bool use_rw)
{
    const int32_t b0 = h->begin, e0 = h->end;
    int64_t len = ((int64_t)(e0 - b0) + kLastChunks - 1) / kLastChunks;
    len = (len + kTile - 1) / kTile * kTile;
    int rc = CF_OK;
    for (int c = 0; c < kLastChunks && rc == CF_OK; ++c) {
        const int64_t cb = (int64_t)b0 + c * len, ce = cb + len < e0 ? cb + len : e0;
        if (cb >= e0) break;
        h->begin = (int32_t)cb;
        h->end = (int32_t)ce;
        h->tree_skip_build = c > 0;
        rc = launch_fit(h, use_rw);
        if (rc == CF_OK) rc = launch_residuals(h);
        if (rc != CF_OK) break;
        const size_t bytes = sizeof(double) * (size_t)(ce - cb);
        cudaError_t e = cudaEventRecord(h->ev_copy[0], h->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(h->early_ys + cb, h->ys + cb, bytes, cudaMemcpyDeviceToHost, h->copy_stream);
        if (e == cudaSuccess && h->early_res)
            e = cudaMemcpyAsync(h->early_res + cb, h->res + cb, bytes, cudaMemcpyDeviceToHost, h->copy_stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "chunked last pass", __FILE__, __LINE__);
    }
    h->begin = b0;
    h->end = e0;
    h->tree_skip_build = false;
    return rc;
}

// `passes` fit passes; the first uses rw iff first_uses_rw (lowess :657-714)
static int32_t run_passes(cf_lowess* h, int32_t passes, bool first_uses_rw, cf_stats* st)
{
    if (h->nranks > 1 && !h->peer_attached && passes > 1) {
        // the median and the weights of a robustness step need every shard: without attached
        // peers (or a host runtime driving the stage calls) later passes would be silently wrong
        set_error("handle holds shard %d of %d but no peers are attached: only single passes can run",
                  (int)h->rank, (int)h->nranks);
        return CF_INVALID_INPUT_ERROR;
    }
    CF_TRY(select_prepare(h));
    CF_TRY(peer_clear_error(h));
    std::vector<cudaEvent_t> evs;
    const bool timed = st != nullptr;
    auto mark = [&]() -> int {
        if (!timed) return CF_OK;
        cudaEvent_t e;
        CF_CUDA(cudaEventCreate(&e));
        evs.push_back(e);
        CF_CUDA(cudaEventRecord(e, h->stream));
        return CF_OK;
    };
    int rc = CF_OK;
    bool use_rw = first_uses_rw;
    bool early_rw_issued = false;
    h->early_rw_missed = nullptr;
    h->timing = timed;
    NvtxRange whole("cf_lowess passes");
    for (int32_t p = 0; p < passes && rc == CF_OK; ++p) {          // :657
        if ((rc = mark()) != CF_OK) break;
        const bool skip_fit = p == 0 && h->first_fit_done;         // issued chunk by chunk during the upload
        const bool chunked = p + 1 == passes && !skip_fit && h->early_ys && h->copy_stream && method_is_tree(h) &&
                             h->nseries == 1 && !(h->delta > 0.0) && h->end - h->begin >= (1 << 22);
        if (chunked) {
            NvtxRange r("lowest pass (last, in pieces beside the download)");
            if ((rc = last_pass_chunked(h, use_rw)) != CF_OK) break;
            h->last_in_pieces = true;
            if ((rc = mark()) != CF_OK) break;
            break;
        }
        if (!skip_fit) {
            NvtxRange r(use_rw ? "lowest pass (robustness weights)" : "lowest pass (first)");
            if ((rc = launch_fit(h, use_rw)) != CF_OK) break;      // :662-693
        }
        if ((rc = mark()) != CF_OK) break;
        NvtxRange r2(p + 1 < passes ? "residuals + median + bisquare" : "residuals");
        if (p + 1 < passes) {                                      // :696
            if ((rc = residuals_and_robust_update(h)) != CF_OK) break;   // :695, :697-713
            use_rw = true;
            if (p + 2 == passes && h->early_rw && h->copy_stream) {
                // m_weights are final: their download runs beside the last fit pass
                cudaError_t e = cudaEventRecord(h->ev_copy[0], h->stream);
                if (e == cudaSuccess) e = cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0);
                if (e == cudaSuccess && h->end > h->begin)
                    e = cudaMemcpyAsync(h->early_rw + h->begin, h->rw + h->begin,
                                        sizeof(double) * (size_t)(h->end - h->begin), cudaMemcpyDeviceToHost,
                                        h->copy_stream);
                if (e != cudaSuccess) { rc = cuda_fail(e, "early download of the robustness weights", __FILE__, __LINE__); break; }
                early_rw_issued = true;
            }
        } else {
            if ((rc = launch_residuals(h)) != CF_OK) break;        // :695
        }
    }
    if (rc == CF_OK) rc = mark();
    if (rc == CF_OK) {
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "lowess passes", __FILE__, __LINE__);
    }
    if (rc == CF_OK) rc = peer_check(h);
    if (rc == CF_OK && timed) {
        // events: [fit0 start, fit0 end, fit1 start, fit1 end, ..., final]
        float fit = 0.f, rob = 0.f, ms = 0.f;
        for (size_t k = 0; k + 1 < evs.size(); ++k) {
            cudaEventElapsedTime(&ms, evs[k], evs[k + 1]);
            if (k % 2 == 0) fit += ms; else rob += ms;
        }
        st->ms_fit = fit;
        st->ms_robust = rob;
        float build = 0.f;
        for (size_t k = 0; k + 1 < h->marks.size(); k += 2) {
            cudaEventElapsedTime(&ms, h->marks[k], h->marks[k + 1]);
            build += ms;
        }
        st->ms_tree_build = build;
    }
    h->timing = false;
    // asked for and there are final weights (a robustness update ran), but the copy was not issued
    // (no copy stream): the fetch that follows must not assume the weights have left
    if (rc == CF_OK && passes >= 2 && h->early_rw && !early_rw_issued) h->early_rw_missed = h->early_rw;
    h->early_rw = nullptr;                 // one-shot
    h->early_ys = h->early_res = nullptr;
    h->first_fit_done = false;
    for (cudaEvent_t e : h->marks) cudaEventDestroy(e);
    h->marks.clear();
    for (cudaEvent_t e : evs) cudaEventDestroy(e);
    return rc;
}

int32_t cf_lowess_run(cf_lowess* h, double f, int32_t nsteps, cf_stats* st)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (nsteps < 0) { set_error("nsteps must be >= 0"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    if (st) memset(st, 0, sizeof(*st));
    const int l0 = h->launches, f0 = h->fit_launches;
    CF_CUDA(cudaEventRecord(h->ev[0], h->stream));
    CF_TRY(begin_stage(h, f));
    if (h->n >= 2) CF_TRY(launch_peer_spans(h));    // attached handles: all ranks call run together
    CF_CUDA(cudaEventRecord(h->ev[1], h->stream));
    if (h->n < 2) {                                                 // :651-654
        if (h->n == 1) CF_CUDA(cudaMemcpyAsync(h->ys, h->y, sizeof(double) * (size_t)h->nseries, cudaMemcpyDeviceToDevice, h->stream));
        CF_CUDA(cudaStreamSynchronize(h->stream));
        h->early_rw = h->early_rw_missed = nullptr;                 // one-shot; there are no weights
    } else {
        CF_TRY(run_passes(h, nsteps + 1, false, st));
    }
    CF_CUDA(cudaEventRecord(h->ev[2], h->stream));
    CF_CUDA(cudaEventSynchronize(h->ev[2]));
    if (st) {
        cudaEventElapsedTime(&st->ms_total, h->ev[0], h->ev[2]);
        cudaEventElapsedTime(&st->ms_windows, h->ev[0], h->ev[1]);
        st->launches = h->launches - l0;
        st->fit_launches = h->fit_launches - f0;
        st->ns = h->ns;
        st->passes = nsteps + 1;
        if (nsteps > 0 && h->n >= 2) {
            SelectState s;
            CF_CUDA(cudaMemcpy(&s, h->sel, sizeof(s), cudaMemcpyDeviceToHost));
            st->cmad_last = s.cmad;
        }
    }
    return CF_OK;
}

int32_t cf_lowess_run_from_rw(cf_lowess* h, double f, const double* rw_in, int32_t passes)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (passes < 1) { set_error("passes must be >= 1"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    CF_TRY(begin_stage(h, f));
    if (h->n < 2) { set_error("n < 2"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(launch_peer_spans(h));
    if (rw_in)
        CF_CUDA(cudaMemcpyAsync(h->rw, rw_in, sizeof(double) * (size_t)h->n, cudaMemcpyHostToDevice, h->stream));
    return run_passes(h, passes, rw_in != nullptr, nullptr);
}

// One pass of the DIRECT kernel over a strided sample of its output tiles (see the header).
int32_t cf_lowess_fit_sample(cf_lowess* h, double f, int32_t use_rw, int32_t first_tile,
                             int32_t tile_stride, int32_t ntiles, float* ms)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    if (first_tile < 0 || tile_stride < 1 || ntiles < 1) {
        set_error("bad tile sample (first %d, stride %d, count %d)", (int)first_tile, (int)tile_stride, (int)ntiles);
        return CF_INVALID_INPUT_ERROR;
    }
    if (h->nseries != 1 || h->delta > 0.0) {
        set_error("cf_lowess_fit_sample: single series with delta = 0 only");
        return CF_INVALID_INPUT_ERROR;
    }
    CF_TRY(use_device(h));
    CF_TRY(begin_stage(h, f));
    if (h->n < 2) { set_error("n < 2"); return CF_INVALID_INPUT_ERROR; }
    const int method = h->method;
    h->method = CF_METHOD_DIRECT;
    h->tile_first = first_tile; h->tile_stride = tile_stride; h->tile_count = ntiles; h->tile_ksplit = 1;
    cudaEventRecord(h->ev[3], h->stream);
    int rc = launch_fit(h, use_rw != 0);
    cudaEventRecord(h->ev[4], h->stream);
    h->tile_count = 0;
    h->method = method;
    if (rc != CF_OK) return rc;
    CF_CUDA(cudaEventSynchronize(h->ev[4]));
    if (ms) CF_CUDA(cudaEventElapsedTime(ms, h->ev[3], h->ev[4]));
    return CF_OK;
}

int32_t cf_lowess_fetch(cf_lowess* h, double* ys, double* rw, double* res)
{
    if (!h || !h->init) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    CF_TRY(use_device(h));
    const size_t bytes = sizeof(double) * (size_t)h->n * (size_t)h->nseries;
    if (ys) CF_CUDA(cudaMemcpyAsync(ys, h->ys, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (rw) CF_CUDA(cudaMemcpyAsync(rw, h->rw, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (res) CF_CUDA(cudaMemcpyAsync(res, h->res, bytes, cudaMemcpyDeviceToHost, h->stream));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    return CF_OK;
}

int32_t cf_lowess_fetch_xy(cf_lowess* h, double* xs, double* ysrt)
{
    if (!h || !h->init) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    CF_TRY(use_device(h));
    CF_TRY(wait_pending_y(h));
    const size_t bytes = sizeof(double) * (size_t)h->n * (size_t)h->nseries;
    if (xs) CF_CUDA(cudaMemcpyAsync(xs, h->x, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (ysrt) CF_CUDA(cudaMemcpyAsync(ysrt, h->y, bytes, cudaMemcpyDeviceToHost, h->stream));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    return CF_OK;
}

__global__ void plus_one_kernel(const int32_t* __restrict__ in, int32_t* __restrict__ out,
                                int32_t n, int32_t add)
{
    const int32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[i] + add;
}

int32_t cf_lowess_windows(cf_lowess* h, double f, int32_t* nleft, int32_t* nright, int32_t* lo9,
                          int32_t* hi9, int32_t* lo1, int32_t* hi1, int32_t* leader)
{
    if (!h) { set_error("null handle"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(use_device(h));
    CF_TRY(begin_stage(h, f));
    if (h->n < 2) return CF_OK;
    const int32_t n = h->n * h->nseries;     // all series (indices stay per series)
    int32_t* tmp = nullptr;
    CF_CUDA(cudaMalloc(&tmp, sizeof(int32_t) * (size_t)n));
    struct { const int32_t* src; int32_t* dst; int32_t add; } jobs[] = {
        {h->win.nleft, nleft, 1}, {h->win.nleft, nright, h->ns},   // nright = nleft + ns - 1
        {h->win.lo9, lo9, 1},     {h->win.hi9, hi9, 1},
        {h->win.lo1, lo1, 1},     {h->win.hi1, hi1, 1},
        {h->win.leader, leader, 1},
    };
    int rc = CF_OK;
    for (auto& j : jobs) {
        if (!j.dst) continue;
        plus_one_kernel<<<(n + 255) / 256, 256, 0, h->stream>>>(j.src, tmp, n, j.add);
        cudaError_t e = cudaMemcpyAsync(j.dst, tmp, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { rc = cuda_fail(e, "windows fetch", __FILE__, __LINE__); break; }
    }
    cudaFree(tmp);
    return rc;
}

// ---- stage API -----------------------------------------------------------------
int32_t cf_lowess_stage_begin(cf_lowess* h, double f)
{
    if (!h) return CF_INVALID_INPUT_ERROR;
    CF_TRY(use_device(h));
    CF_TRY(begin_stage(h, f));
    if (h->n < 2) { set_error("n < 2"); return CF_INVALID_INPUT_ERROR; }
    CF_TRY(select_prepare(h));
    return launch_peer_spans(h);         // attached: every rank learns whose windows reach what
}
int32_t cf_lowess_stage_fit(cf_lowess* h, int32_t use_rw)
{
    if (!h || !h->init) return CF_NO_DATA_DEFINED_ERROR;
    CF_TRY(use_device(h));
    CF_TRY(launch_fit(h, use_rw != 0));
    return launch_residuals(h);
}
// round 0: the B0-bit histogram of |res| over the shard (left in CF_BUF_HIST0 for the host
// collective); round k >= 1: the k-th W-bit digit of the shard's candidates (CF_BUF_HIST)
int32_t cf_lowess_stage_hist(cf_lowess* h, int32_t round)
{
    if (!h || !h->init || round < 0 || round > select_rounds(h)) return CF_INVALID_INPUT_ERROR;
    CF_TRY(use_device(h));
    return round == 0 ? launch_hist0(h, false, true) : launch_cand_round(h, round, true);
}
// after the counters were summed over the ranks: round 0 fixes the bucket and collects the
// shard's candidates; round k >= 1 extends the prefix (the last one forms cmad, c1, c9)
int32_t cf_lowess_stage_pick(cf_lowess* h, int32_t round)
{
    if (!h || !h->init || round < 0 || round > select_rounds(h)) return CF_INVALID_INPUT_ERROR;
    CF_TRY(use_device(h));
    if (round > 0) return launch_cand_pick(h, round);
    CF_TRY(launch_pick0(h));
    const bool attached = h->peer_attached;      // host-driven protocol: candidates stay local
    h->peer_attached = false;
    const int rc = launch_compact(h);
    h->peer_attached = attached;
    return rc;
}
int32_t cf_lowess_stage_weights(cf_lowess* h)
{
    if (!h || !h->init) return CF_NO_DATA_DEFINED_ERROR;
    CF_TRY(use_device(h));
    return launch_bisquare(h);
}
int32_t cf_lowess_sync(cf_lowess* h)
{
    if (!h) return CF_INVALID_INPUT_ERROR;
    CF_TRY(use_device(h));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    return CF_OK;
}

void* cf_lowess_buffer(cf_lowess* h, int32_t which)
{
    if (!h) return nullptr;
    switch (which) {
        case CF_BUF_X: return h->x;
        case CF_BUF_Y: return h->y;
        case CF_BUF_YS: return h->ys;
        case CF_BUF_RW: return h->rw;
        case CF_BUF_RES: return h->res;
        case CF_BUF_HIST: return h->hist;
        case CF_BUF_HIST0: return h->hist0;
        case CF_BUF_CMAD: return &h->sel->cmad;
        default: return nullptr;
    }
}

int32_t cf_fp64_peak(int32_t device, double ms_target, double* tflops, double* sm_mhz_est)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device available");
        return CF_DEVICE_ERROR;
    }
    return fp64_peak(device, ms_target, tflops, sm_mhz_est);
}

// ---- stateless host-buffer entry points -------------------------------------------
// The stateless calls keep one idle handle per device alive between calls so that a
// caller looping over smooth() (the Fortran type does exactly that) pays for device
// allocations, stream and event creation once, not per call.
static std::mutex g_pool_mu;
static std::vector<cf_lowess*> g_pool;

static int32_t pool_acquire(cf_lowess** out)
{
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess) {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (size_t k = 0; k < g_pool.size(); ++k) {
            if (g_pool[k]->device == dev) {
                *out = g_pool[k];
                g_pool.erase(g_pool.begin() + (long)k);
                (*out)->method = stateless_method();
                return CF_OK;
            }
        }
    }
    CF_TRY(cf_lowess_create(out, -1, nullptr));
    (*out)->method = stateless_method();
    return CF_OK;
}

static void pool_release(cf_lowess* h, int rc)
{
    if (!h) return;
    if (rc == CF_OK) {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        if (g_pool.size() < 8) {
            h->rank = 0; h->nranks = 1;
            h->delta = 0.0;
            g_pool.push_back(h);
            return;
        }
    }
    cf_lowess_destroy(h);
}

int32_t cf_sort_xy(const double* x, const double* y, int32_t n, double* x_sorted,
                   double* y_sorted, int32_t* perm)
{
    if (n < 0 || (n > 0 && (!x || !y || !x_sorted || !y_sorted))) {
        set_error("Input array sizes must match.");
        return CF_ARRAY_SIZE_ERROR;
    }
    if (n == 0) return CF_OK;
    cf_lowess* h = nullptr;
    CF_TRY(pool_acquire(&h));
    int rc = CF_OK;
    double *dx = nullptr, *dy = nullptr, *ox = nullptr, *oy = nullptr;
    int32_t* dp = nullptr;
    const size_t bytes = sizeof(double) * (size_t)n;
    cudaError_t e = cudaMalloc(&dx, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&dy, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&ox, bytes);
    if (e == cudaSuccess) e = cudaMalloc(&oy, bytes);
    if (e == cudaSuccess && perm) e = cudaMalloc(&dp, sizeof(int32_t) * (size_t)n);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dx, x, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dy, y, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e != cudaSuccess) rc = cuda_fail(e, "cf_sort_xy upload", __FILE__, __LINE__);
    if (rc == CF_OK) rc = sort_xy_device(dx, dy, n, ox, oy, dp, h->stream, &h->launches);
    if (rc == CF_OK) {
        e = cudaMemcpyAsync(x_sorted, ox, bytes, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(y_sorted, oy, bytes, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess && perm) e = cudaMemcpyAsync(perm, dp, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "cf_sort_xy download", __FILE__, __LINE__);
    }
    cudaFree(dx); cudaFree(dy); cudaFree(ox); cudaFree(oy); cudaFree(dp);
    pool_release(h, rc);
    return rc;
}

__global__ void __launch_bounds__(256)
fill_kernel(double* __restrict__ x, int64_t n, double v)
{
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) x[k] = v;
}

__global__ void set_build_span_kernel(int32_t* sp, int32_t lo, int32_t hi, int32_t top_hi)
{
    sp[0] = lo; sp[1] = hi; sp[2] = 0; sp[3] = top_hi;
}

// cf_lowess_smooth on a long series, tree evaluator: the series is uploaded in chunks and the FIRST
// pass (windows, tree build, evaluation) follows the upload chunk by chunk instead of waiting for
// all of it (at n = 1e8 the upload is 32 ms, the first pass 25 ms).
//  * x starts as +inf everywhere except x(n) (lowest() needs x(n) - x(1), :535), so every guided
//    search of windows.cu that probes past the arrived elements meets a monotone predicate;
//  * a point is processed once x(i + ns) has arrived (windows.cu reads x(L + ns), L <= i; its
//    window ends at or before i + ns - 1) and the run of values equal to its window's last one
//    has ended;
//  * chunks are multiples of every power-of-two node span the chunk-gated build kernels handle
//    (binary levels with more than 256 nodes: < n / 256 elements; chain blocks: <= ns / 2), so
//    no such node straddles a chunk; the few top levels are rebuilt over everything that has arrived.
// Node values, windows and fitted values are those of the undivided pass, bit for bit.
static int first_pass_pipelined(cf_lowess* h, const double* x, const double* y, double f)
{
    const int32_t n = h->n;
    const int32_t ns = cf_lowess_span(n, f);
    int64_t align = 1;
    while (align < (int64_t)n / 256 || align < (int64_t)ns / 2) align <<= 1;
    if (align < 16 * 512) align = 16 * 512;
    int64_t chunk = ((int64_t)n / 12 + align - 1) / align * align;
    if (chunk < align) chunk = align;
    const int C = (int)(((int64_t)n + chunk - 1) / chunk);
    while ((int)h->chunk_events.size() < C) {
        cudaEvent_t e;
        CF_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        h->chunk_events.push_back(e);
    }
    if (!h->build_span_buf) CF_CUDA(cudaMalloc(&h->build_span_buf, 4 * sizeof(int32_t)));
    int32_t* const span_dev = h->build_span_buf;
    h->ns = ns;
    CF_TRY(select_prepare(h));
    fill_kernel<<<296, 256, 0, h->stream>>>(h->x, (int64_t)n - 1, (double)INFINITY);
    CF_CUDA(cudaGetLastError());
    CF_CUDA(cudaMemcpyAsync(h->x + (n - 1), x + (n - 1), sizeof(double), cudaMemcpyHostToDevice, h->stream));
    CF_CUDA(cudaEventRecord(h->ev_copy[0], h->stream));
    CF_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0));
    h->launches++;
    int rc = CF_OK;
    int32_t p0 = 0;
    bool first_windows = true;
    for (int k = 0; k < C && rc == CF_OK; ++k) {
        const int64_t e0 = (int64_t)k * chunk, e1 = e0 + chunk < n ? e0 + chunk : n;
        const size_t bytes = sizeof(double) * (size_t)(e1 - e0);
        cudaError_t e = cudaMemcpyAsync(h->x + e0, x + e0, bytes, cudaMemcpyHostToDevice, h->copy_stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(h->y + e0, y + e0, bytes, cudaMemcpyHostToDevice, h->copy_stream);
        if (e == cudaSuccess) e = cudaEventRecord(h->chunk_events[(size_t)k], h->copy_stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(h->stream, h->chunk_events[(size_t)k], 0);
        if (e != cudaSuccess) { rc = cuda_fail(e, "pipelined upload", __FILE__, __LINE__); break; }
        // points whose windows lie inside what has arrived: i + ns < e1, and -- a window of equal
        // abscissas (h = 0) takes every element equal to xs, however far the run goes (:546) -- the
        // window's last value must be below the last arrived one, whose run may continue in the next chunk
        int64_t p1 = n;
        if (k < C - 1) {
            const int64_t tfirst = std::lower_bound(x, x + e1, x[e1 - 1]) - x;    // first element of that run
            p1 = tfirst - ns + 1;
        }
        if (p1 < p0) p1 = p0;
        if (p1 > p0) {
            rc = launch_windows_range(h, ns, p0, (int32_t)p1, first_windows);
            first_windows = false;
            if (rc != CF_OK) break;
        }
        // the tree over the new elements (+ the top levels over everything so far)
        set_build_span_kernel<<<1, 1, 0, h->stream>>>(span_dev, (int32_t)e0, (int32_t)(e1 - 1), (int32_t)(e1 - 1));
        h->build_span = span_dev;
        h->tree_build_only = true;
        h->begin = 0; h->end = n;
        rc = launch_fit(h, false);
        h->tree_build_only = false;
        h->build_span = nullptr;
        if (rc != CF_OK) break;
        if (p1 > p0) {
            h->begin = p0; h->end = (int32_t)p1;
            h->tree_skip_build = true;
            rc = launch_fit(h, false);                 // :662-693 for these points, first pass (userw = .false.)
            h->tree_skip_build = false;
            p0 = (int32_t)p1;
        }
    }
    h->begin = 0; h->end = n;
    if (rc == CF_OK) {
        h->win_f = f;
        h->first_fit_done = true;
    }
    return rc;
}

int32_t cf_lowess_smooth(const double* x_sorted, const double* y, int32_t n, double f,
                         int32_t nsteps, double* ys, double* rw, double* res)
{
    if (n < 0 || (n > 0 && (!x_sorted || !y || !ys))) {
        set_error("Input array sizes must match.");
        return CF_ARRAY_SIZE_ERROR;
    }
    if (nsteps < 0) { set_error("nsteps must be >= 0"); return CF_INVALID_INPUT_ERROR; }
    cf_lowess* h = nullptr;
    CF_TRY(pool_acquire(&h));
    int rc = CF_OK;
    // long series: the copies are worth overlapping with the kernels (a few ms per 100 MB)
    const bool overlap = n >= (1 << 20) && (f == f);
    if (!overlap) {
        rc = cf_lowess_init(h, x_sorted, y, n, 0);
        if (rc == CF_OK) rc = cf_lowess_run(h, f, nsteps, nullptr);
        if (rc == CF_OK) {
            if (n < 2) rc = cf_lowess_fetch(h, ys, nullptr, nullptr);     // rw, res untouched (:651-654)
            else rc = cf_lowess_fetch(h, ys, nsteps > 0 ? rw : nullptr, res);
        }
        pool_release(h, rc);
        return rc;
    }
    auto body = [&]() -> int {
        CF_TRY(layout_series(h, n));                                  // ls_init, srt = .false. (:800-803)
        CF_TRY(ensure_copy_stream(h));
        const size_t bytes = sizeof(double) * (size_t)n;
        // tree evaluator: the first pass follows the upload chunk by chunk, and ys / res leave piece
        // by piece during the last pass (run_passes)
        h->ns = cf_lowess_span(n, f);
        const bool pieces = method_is_tree(h) && n >= (1 << 22);
        if (pieces) {
            CF_TRY(first_pass_pipelined(h, x_sorted, y, f));
        } else {
            CF_CUDA(cudaMemcpyAsync(h->x, x_sorted, bytes, cudaMemcpyHostToDevice, h->stream));
            CF_CUDA(cudaEventRecord(h->ev_copy[0], h->stream));
            // y follows x over the bus on the second stream while the window search (x only) runs
            CF_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_copy[0], 0));
            CF_CUDA(cudaMemcpyAsync(h->y, y, bytes, cudaMemcpyHostToDevice, h->copy_stream));
            CF_CUDA(cudaEventRecord(h->ev_copy[1], h->copy_stream));
            h->win_f = -1.0;
            CF_TRY(begin_stage(h, f));                                // windows
            CF_CUDA(cudaStreamWaitEvent(h->stream, h->ev_copy[1], 0));
        }
        h->early_rw = (rw && nsteps > 0) ? rw : nullptr;
        h->early_ys = pieces ? ys : nullptr;
        h->early_res = pieces ? res : nullptr;
        h->last_in_pieces = false;
        int rc2 = cf_lowess_run(h, f, nsteps, nullptr);
        h->early_rw = nullptr;
        h->early_ys = h->early_res = nullptr;
        if (rc2 != CF_OK) { cudaStreamSynchronize(h->copy_stream); return rc2; }
        if (h->early_rw_missed) {
            CF_CUDA(cudaMemcpyAsync(rw, h->rw, bytes, cudaMemcpyDeviceToHost, h->stream));
            h->early_rw_missed = nullptr;
        }
        if (!h->last_in_pieces) {
            CF_CUDA(cudaMemcpyAsync(ys, h->ys, bytes, cudaMemcpyDeviceToHost, h->stream));
            if (res) CF_CUDA(cudaMemcpyAsync(res, h->res, bytes, cudaMemcpyDeviceToHost, h->stream));
        }
        CF_CUDA(cudaStreamSynchronize(h->stream));
        CF_CUDA(cudaStreamSynchronize(h->copy_stream));
        return CF_OK;
    };
    rc = body();
    if (rc != CF_OK && h->copy_stream) cudaStreamSynchronize(h->copy_stream);
    pool_release(h, rc);
    return rc;
}

int32_t cf_lowess_smooth_ex(const double* x_sorted, const double* y, int32_t n, double f, int32_t nsteps,
                            int32_t method, double* ys, double* rw, double* res)
{
    if (method != CF_METHOD_DIRECT && method != CF_METHOD_TREE && method != CF_METHOD_AUTO) {
        set_error("method must be CF_METHOD_DIRECT, CF_METHOD_TREE or CF_METHOD_AUTO");
        return CF_INVALID_INPUT_ERROR;
    }
    const int before = t_call_method;
    t_call_method = method;                    // seen by pool_acquire of this thread only
    const int32_t rc = cf_lowess_smooth(x_sorted, y, n, f, nsteps, ys, rw, res);
    t_call_method = before;
    return rc;
}

int32_t cf_lowess_smooth_delta(const double* x_sorted, const double* y, int32_t n, double f,
                               int32_t nsteps, double delta, double* ys, double* rw, double* res)
{
    if (n < 0 || (n > 0 && (!x_sorted || !y || !ys))) {
        set_error("Input array sizes must match.");
        return CF_ARRAY_SIZE_ERROR;
    }
    if (nsteps < 0) { set_error("nsteps must be >= 0"); return CF_INVALID_INPUT_ERROR; }
    cf_lowess* h = nullptr;
    CF_TRY(pool_acquire(&h));
    int rc = cf_lowess_set_delta(h, delta);
    if (rc == CF_OK) rc = cf_lowess_init(h, x_sorted, y, n, 0);
    if (rc == CF_OK) rc = cf_lowess_run(h, f, nsteps, nullptr);
    if (rc == CF_OK) {
        if (n < 2) rc = cf_lowess_fetch(h, ys, nullptr, nullptr);
        else rc = cf_lowess_fetch(h, ys, nsteps > 0 ? rw : nullptr, res);
    }
    pool_release(h, rc);
    return rc;
}

int32_t cf_lowess_smooth_batched(const double* x_sorted, const double* y, int32_t nseries,
                                 int32_t n, double f, int32_t nsteps, double* ys, double* rw,
                                 double* res)
{
    if (nseries < 0 || n < 0) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    if (nsteps < 0) { set_error("nsteps must be >= 0"); return CF_INVALID_INPUT_ERROR; }
    if (nseries == 0 || n == 0) return CF_OK;
    if (!x_sorted || !y || !ys) { set_error("Input array sizes must match."); return CF_ARRAY_SIZE_ERROR; }
    // Independent series are a grid dimension of every kernel (blockIdx.y), so one
    // launch covers a whole group of series; groups are bounded to ~2 GB of device
    // state and alternate between two handles (streams) so that the copies of one
    // group overlap the kernels of the other.
    int64_t per = (int64_t)(2.0e9 / (64.0 * (double)n));
    if (per < 1) per = 1;
    if (per > 32768) per = 32768;            // gridDim.y limit is 65535
    if (per > nseries) per = nseries;
    cf_lowess* hs[2] = {nullptr, nullptr};
    int rc = cf_lowess_create(&hs[0], -1, nullptr);
    if (rc == CF_OK && per < nseries) rc = cf_lowess_create(&hs[1], -1, nullptr);
    auto enqueue = [&](cf_lowess* h, int32_t s0, int32_t cnt) -> int {
        const size_t off = (size_t)s0 * (size_t)n, bytes = sizeof(double) * (size_t)n * (size_t)cnt;
        h->init = false;
        h->win_f = -1.0;
        h->rank = 0; h->nranks = 1;
        h->x_partial = false;
        CF_TRY(reserve(h, n, cnt));
        h->n = n;
        h->nseries = cnt;
        compute_shard(h);
        CF_TRY(reserve_rw(h));
        CF_CUDA(cudaMemcpyAsync(h->x, x_sorted + off, bytes, cudaMemcpyHostToDevice, h->stream));
        CF_CUDA(cudaMemcpyAsync(h->y, y + off, bytes, cudaMemcpyHostToDevice, h->stream));
        h->init = true;
        CF_TRY(begin_stage(h, f));
        if (n < 2) {
            CF_CUDA(cudaMemcpyAsync(h->ys, h->y, bytes, cudaMemcpyDeviceToDevice, h->stream));
        } else {
            bool use_rw = false;
            CF_TRY(select_prepare(h));
            for (int32_t p = 0; p <= nsteps; ++p) {
                CF_TRY(launch_fit(h, use_rw));
                if (p < nsteps) { CF_TRY(residuals_and_robust_update(h)); use_rw = true; }
                else CF_TRY(launch_residuals(h));
            }
        }
        CF_CUDA(cudaMemcpyAsync(ys + off, h->ys, bytes, cudaMemcpyDeviceToHost, h->stream));
        if (n >= 2) {
            if (rw && nsteps > 0) CF_CUDA(cudaMemcpyAsync(rw + off, h->rw, bytes, cudaMemcpyDeviceToHost, h->stream));
            if (res) CF_CUDA(cudaMemcpyAsync(res + off, h->res, bytes, cudaMemcpyDeviceToHost, h->stream));
        }
        return CF_OK;
    };
    int which = 0;
    for (int64_t s0 = 0; s0 < nseries && rc == CF_OK; s0 += per, which ^= 1) {
        cf_lowess* h = hs[which] ? hs[which] : hs[0];
        // the handle's previous group must have left the device before its buffers are reused
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) { rc = cuda_fail(e, "batched smooth", __FILE__, __LINE__); break; }
        const int32_t cnt = (int32_t)((nseries - s0) < per ? (nseries - s0) : per);
        rc = enqueue(h, (int32_t)s0, cnt);
    }
    for (cf_lowess* h : hs) {
        if (!h) continue;
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess && rc == CF_OK) rc = cuda_fail(e, "batched smooth", __FILE__, __LINE__);
        cf_lowess_destroy(h);
    }
    return rc;
}

}  // extern "C"
