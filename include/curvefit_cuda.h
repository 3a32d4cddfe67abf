/*
 * curvefit_cuda.h -- C ABI of libcurvefit_cuda (sm_100a only, no CPU fallback).
 *
 * Drop-in boundary for ONE path of jchristopherson/curvefit: robust LOWESS
 * smoothing, i.e. lowess_smoothing%initialize / %smooth
 * (src/curvefit_regression.f90:218-247 type, :740-811 ls_init, :832-868
 * ls_smooth, :625-715 lowess, :513-586 lowest).  The reference has no FFI layer
 * for this path (its only C interop is FERROR's, ferror.h); these entry points
 * are what a `bind(C)` interface block in curvefit_regression.f90 would bind
 * (see fortran/curvefit_regression_cuda.f90 and INTEGRATION.md).
 *
 * Conventions
 *  - plain pointers and sizes, no C++/torch types; all arrays real64 / int32
 *    like the reference (real(real64), integer(int32));
 *  - every function returns 0 on success, else one of the CF_* codes below (the
 *    reference's own values) -- it never exits, prints or throws;
 *    cf_last_error() gives a thread-local detail string;
 *  - "host" pointers are caller-owned and never retained past the call;
 *  - indices written to int32 outputs are 1-BASED like the reference's
 *    nleft/nright (src/curvefit_regression.f90:658-669);
 *  - one handle must not be used from two threads at once (the reference's
 *    object is intent(inout) in smooth, :834); distinct handles are independent.
 */
#ifndef CURVEFIT_CUDA_H_
#define CURVEFIT_CUDA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define CF_API __declspec(dllexport)
#else
#define CF_API __attribute__((visibility("default")))
#endif

/* Error codes.  202/105/300/201 are the reference's (src/curvefit_core.f90:34-40
 * via nonlin_constants.f90:18,20 and linalg_constants.f90:23); 303 is new
 * (300-302 are taken, curvefit_core.f90:38,43,54) and means a CUDA failure. */
#define CF_OK 0
#define CF_OUT_OF_MEMORY_ERROR 105
#define CF_INVALID_INPUT_ERROR 201
#define CF_ARRAY_SIZE_ERROR 202
#define CF_NO_DATA_DEFINED_ERROR 300
#define CF_DEVICE_ERROR 303

/* Library version (major*10000 + minor*100 + patch). */
CF_API int32_t cf_version(void);

/* Copies the calling thread's last error detail (NUL terminated) into buf;
 * returns the length that was needed. */
CF_API int32_t cf_last_error(char* buf, int32_t len);

/* Span of the local fit: ns = max(min(int(f*n), n), 2)
 * -- src/curvefit_regression.f90:647 (truncation, not ceil). */
CF_API int32_t cf_lowess_span(int32_t n, double f);

/* ------------------------------------------------------------------------
 * Stateless entry points (HOST buffers) -- what the Fortran shim calls.
 * ---------------------------------------------------------------------- */

/* ls_init with srt=.true., data part (src/curvefit_regression.f90:791-799):
 * x_sorted = ascending x, y_sorted = y co-permuted.  Replaces
 * sort(m_x, indices, .true.) (LINALG linalg_sorting.f90:38-78, 277-354) with a
 * stable device LSD radix sort; `perm` (may be NULL) receives the 1-based
 * tracking indices.  In/out arrays may not alias. */
CF_API int32_t cf_sort_xy(const double* x, const double* y, int32_t n,
                          double* x_sorted, double* y_sorted, int32_t* perm);

/* lowess(x, y, f, nsteps, delta=0, ys, rw, res)
 * (src/curvefit_regression.f90:625-715, as called from ls_smooth :866 with
 * nsteps = 2).  x ascending, all arrays length n.  n < 2: ys = y, rw/res
 * untouched (:651-654).  nsteps = 0: rw untouched (never written, :696). */
CF_API int32_t cf_lowess_smooth(const double* x_sorted, const double* y, int32_t n,
                                double f, int32_t nsteps,
                                double* ys, double* rw, double* res);

/* cf_lowess_smooth with the evaluator of a lowest() pass given PER CALL (CF_METHOD_DIRECT / _TREE /
 * _AUTO, below) instead of taken from the process-wide default: the choice is local to the calling
 * thread and to this call, so two threads can smooth with different evaluators at the same time and
 * nothing a library deep in the program does with cf_set_default_method changes this call's
 * numerical path.  No counterpart in the reference (it has one evaluator).  Everything else as
 * cf_lowess_smooth.  Rule of thumb: a direct pass costs n * ns pairs at about 1.5e12 pairs/s on a
 * B200 -- beyond n * ns ~ 1e12 choose CF_METHOD_TREE (equal to <= 1e-12 relative). */
CF_API int32_t cf_lowess_smooth_ex(const double* x_sorted, const double* y, int32_t n,
                                   double f, int32_t nsteps, int32_t method,
                                   double* ys, double* rw, double* res);

/* The same with delta > 0 -- lowess's interpolation mode (:674-690): lowest() runs
 * only at points more than delta apart, points in between are linearly interpolated
 * (:675-679), points closer than epsilon to the last fitted one copy it (:685-688).
 * The reference's smooth() always passes delta = 0 (:807, :866), so this is an
 * extension of the drop-in surface; results equal the reference's lowess(delta). */
CF_API int32_t cf_lowess_smooth_delta(const double* x_sorted, const double* y, int32_t n,
                                      double f, int32_t nsteps, double delta,
                                      double* ys, double* rw, double* res);

/* nseries independent series, row-major [nseries][n]; same semantics per row. */
CF_API int32_t cf_lowess_smooth_batched(const double* x_sorted, const double* y,
                                        int32_t nseries, int32_t n, double f,
                                        int32_t nsteps,
                                        double* ys, double* rw, double* res);

/* lowess() of ONE series over several GPUs of one box, from one process (multi.cu) -- the
 * multi-GPU form of cf_lowess_smooth and what the Fortran shim's smooth() binds
 * (src/curvefit_regression.f90:832-868, ls_smooth).  The sorted index range is cut into ndev
 * equal blocks of output points; device k uploads only the elements its block's windows
 * reach (block + the (ns-1)-point halo on either side), the robustness step's exchanges cross
 * NVLink inside the kernels (peer.cu; the devices must be peer-accessible), and every device
 * writes its own block of ys, rw, res into the caller's arrays.  Results are bit-identical to
 * cf_lowess_smooth.  devices = NULL or ndev <= 0: all visible devices (at most 8).  One device,
 * or fewer than 4096 points per device: cf_lowess_smooth on devices[0].  The handles and the
 * peer mapping are kept between calls with the same n and device list;
 * cf_lowess_multi_release() frees them.  Calls are serialised per process.
 * No peer access between the devices (PCIe box without P2P, mixed GPUs, MIG), or no room for the
 * buffers on one of them: the sharded context is freed and the call runs as cf_lowess_smooth on
 * ONE device (the caller's current device if it is in the list, else the first of the list) --
 * same results, never an error a single GPU would not also give.  A device list without peer
 * access is remembered for the life of the process; an out-of-memory one is tried again next call.
 * cf_lowess_multi_active() = number of devices of the cached sharded context (0: none, i.e. the
 * last call ran on one GPU or nothing has run).
 * Memory: every device reserves buffers for all n points (about 60 bytes per point), not only for
 * its block plus halo -- only the copies shrink with the shard -- so the longest series is the one
 * a single GPU holds (n = 1e8: 6 GB per device). */
CF_API int32_t cf_lowess_smooth_multi(const double* x_sorted, const double* y, int32_t n,
                                      double f, int32_t nsteps,
                                      const int32_t* devices, int32_t ndev,
                                      double* ys, double* rw, double* res);
CF_API int32_t cf_lowess_multi_release(void);
CF_API int32_t cf_lowess_multi_active(void);

/* ------------------------------------------------------------------------
 * Neighbours of the LOWESS path (stats.cu).
 * ---------------------------------------------------------------------- */

/* moving_average_1(x, npts) (src/curvefit_regression.f90:427-478), in place on a
 * HOST array: x(i) <- mean of the last npts samples, zeros before the start of
 * the signal (:467-473), mean by the running update of mean_dbl
 * (src/curvefit_statistics.f90:239-242).  Bit-identical to the reference.
 * npts < 2 or npts > n: CF_INVALID_INPUT_ERROR (:452-457), x untouched. */
CF_API int32_t cf_moving_average(double* x, int32_t n, int32_t npts);
/* Same on device memory, out of place, asynchronous on `stream` (a cudaStream_t). */
CF_API int32_t cf_moving_average_device(const void* d_x, void* d_out, int32_t n,
                                        int32_t npts, void* stream);

/* median_dbl (src/curvefit_statistics.f90:258-295) of a host array: an exact radix
 * select of the two middle order statistics instead of the reference's full sort
 * (:286); *z = x(nmidp1) for odd n, half * (x(nmid) + x(nmidp1)) for even n, 0 for
 * n = 0 (:280-283).  x is not modified; a caller that wants the reference's
 * srt = .true. side effect (x sorted on return) calls cf_sort as well. */
CF_API int32_t cf_median(const double* x, int32_t n, double* z);
/* Same on device memory, asynchronous on `stream` (a cudaStream_t; NULL = the legacy
 * default stream, as for cf_moving_average_device); scratch is allocated in stream order. */
CF_API int32_t cf_median_device(const void* d_x, int32_t n, void* d_z, void* stream);

/* cf_median_device for callers that must not allocate device memory.  No counterpart in the
 * reference (it has no device memory at all); it exists because cf_median_device takes its select
 * scratch from cudaMallocAsync, and some callers cannot afford or are not allowed that:
 *   - a process in which another allocator (a framework's caching pool, a fixed arena set up at
 *     start-up) already owns all of the GPU's memory, so a driver allocation fails or evicts;
 *   - devices or virtualised set-ups without stream-ordered memory pools, where cudaMallocAsync
 *     is simply not available;
 *   - latency-bound loops (control, acquisition) that have finished their set-up phase and may not
 *     take the driver's allocation locks or an implicit synchronisation any more.
 * Here the scratch is a fixed array inside the library image: about 8.2 MB of device memory on EVERY
 * GPU that loads libcurvefit_cuda, held until the process exits whether or not this function is
 * ever called.  Limit: n <= CF_MEDIAN_NOALLOC_MAX_N; larger n returns CF_ARRAY_SIZE_ERROR with a
 * message and touches nothing.  Same result as cf_median_device, bit for bit.
 *   - The caller still owns d_x (n doubles) and d_z (one double), both on the CURRENT device, and
 *     `stream` (NULL = legacy default stream) must belong to the current device too.
 *   - The first call on each device looks up the array's device address (one runtime query, no
 *     allocation) and caches it; make that first call outside any stream capture.
 *   - All calls on a device share the one array.  Each call clears the counters it needs itself, so
 *     back-to-back calls on one stream are fine; two calls that can run on the GPU at the same time
 *     (different streams without an ordering between them, or two host threads) race on it and
 *     return garbage.  Serialise them, or use cf_median_device for the concurrent ones.
 * Calls no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine. */
#define CF_MEDIAN_NOALLOC_MAX_N 1048576
CF_API int32_t cf_median_device_noalloc(const void* d_x, int32_t n, void* d_z, void* stream);

/* sort(x, .true.) (LINALG linalg_sorting.f90:12-35 -> DLASRT 'I'): ascending values
 * by the device radix sort.  In and out may not alias. */
CF_API int32_t cf_sort(const double* x, int32_t n, double* x_sorted);

/* ------------------------------------------------------------------------
 * Device-resident handle -- the lowess_smoothing object's state on one GPU
 * (m_x, m_y, m_weights, m_residuals: src/curvefit_regression.f90:218-233).
 * ---------------------------------------------------------------------- */
typedef struct cf_lowess cf_lowess;

typedef struct cf_stats {
    float ms_total;      /* whole run, CUDA events on the handle's stream     */
    float ms_windows;    /* neighbourhood search                              */
    float ms_fit;        /* sum over passes of the local-fit kernel           */
    float ms_robust;     /* residuals + select + bisquare, summed over passes */
    int32_t fit_launches;
    int32_t launches;    /* all kernels of this library launched by the run   */
    int32_t ns;          /* span used                                         */
    int32_t passes;      /* nsteps + 1                                        */
    double cmad_last;    /* cmad of the last robustness iteration (:701)      */
    float ms_tree_build; /* part of ms_fit: per-pass rebuild of the moment tree */
    float reserved;
} cf_stats;

/* device < 0: current device.  stream: a cudaStream_t (as void*) or NULL for a
 * private non-blocking stream owned by the handle. */
CF_API int32_t cf_lowess_create(cf_lowess** h, int32_t device, void* stream);
CF_API void cf_lowess_destroy(cf_lowess* h);

/* ls_init (src/curvefit_regression.f90:740-811): copy host x, y to the device;
 * sort != 0 sorts x ascending and co-permutes y (default of the reference).
 * No monotonicity check when sort == 0, like the reference (:800-803). */
CF_API int32_t cf_lowess_init(cf_lowess* h, const double* x, const double* y,
                              int32_t n, int32_t sort);
/* Same, x and y already in device memory (device-to-device copy). */
CF_API int32_t cf_lowess_init_device(cf_lowess* h, const void* d_x, const void* d_y,
                                     int32_t n, int32_t sort);

/* ls_init with srt = .false. for a range-sharded handle that receives only the elements its
 * shard's windows reach: x_sorted, y are the whole host series (length n), of which elements
 * first .. last (1-based, inclusive) are uploaded; x outside becomes -inf / +inf sentinels
 * (x(1) and x(n) keep their values: lowest() needs x(n) - x(1), :535), y outside is
 * untouched.  cf_lowess_shard_span gives the span the handle's shard (cf_lowess_set_shard)
 * needs for span fraction f: from nleft of the block's first point to nright + 1 of its last one,
 * widened over ties -- at most the (ns-1)-point halo either side, ns / 2 on a grid.  cf_lowess_fetch_block copies only the shard's own block
 * [begin, end) of ys, rw, res into full-length host arrays (each may be NULL). */
CF_API int32_t cf_lowess_shard_span(const cf_lowess* h, const double* x_sorted, int32_t n, double f,
                                    int32_t* first, int32_t* last);
/* The same for an arbitrary block of output points begin .. end (1-based, inclusive); needs no
 * handle and no device: nleft(begin) .. nright(end) + 1 by the reference's own predicate
 * (:664-666), widened over ties. */
CF_API int32_t cf_lowess_block_span(const double* x_sorted, int32_t n, double f, int32_t begin, int32_t end,
                                    int32_t* first, int32_t* last);
CF_API int32_t cf_lowess_init_span(cf_lowess* h, const double* x_sorted, const double* y, int32_t n,
                                   int32_t first, int32_t last);
CF_API int32_t cf_lowess_fetch_block(cf_lowess* h, double* ys, double* rw, double* res);
/* One-shot: during the NEXT cf_lowess_run the shard's block of the final robustness weights
 * (m_weights, final once the last robustness update is done, :704-713) is copied to
 * rw_host[begin .. end) on a second stream while the last fit pass runs; the following
 * cf_lowess_fetch_block (pass rw = NULL there) or cf_lowess_sync completes it.  rw_host is a
 * full-length host array; NULL cancels. */
CF_API int32_t cf_lowess_set_early_weights(cf_lowess* h, double* rw_host);

/* Batched variant: nseries independent series of equal length n, row-major
 * [nseries][n], x already ascending per row (srt = .false. semantics).  Every
 * later call on the handle (run, fetch, windows) then works on all series at
 * once; each series is smoothed exactly as a separate lowess_smoothing object
 * would be.  Sharding (cf_lowess_set_shard) does not apply to batched handles. */
CF_API int32_t cf_lowess_init_batched(cf_lowess* h, const double* x, const double* y,
                                      int32_t nseries, int32_t n);

/* Restrict the fitted output points to block `rank` of `nranks` equal blocks
 * of the sorted index range (multi-GPU range partition).  Default 0 of 1. */
CF_API int32_t cf_lowess_set_shard(cf_lowess* h, int32_t rank, int32_t nranks);
CF_API int32_t cf_lowess_shard(const cf_lowess* h, int32_t* begin, int32_t* end,
                               int32_t* block);

/* How one lowest() pass (:513-586) is evaluated.
 *  CF_METHOD_DIRECT  the O(n * ns) sweep, one (point, element) pair at a time in
 *                    the reference's summation order (fit.cu; FP64-pipe bound);
 *  CF_METHOD_TREE    the O(n log ns) moment-tree evaluator (tree.cu): the same sums
 *                    from per-block power moments, equal to rounding (<= 1e-12
 *                    relative measured; contract 1e-10);
 *  CF_METHOD_AUTO    tree when ns >= 1024, else direct.
 * Default: CF_METHOD_DIRECT.  cf_set_default_method changes, process-wide, the method of
 * handles created afterwards and of the stateless calls (cf_lowess_smooth, _delta,
 * _batched, _multi); there is no environment route. */
#define CF_METHOD_DIRECT 0
#define CF_METHOD_TREE 1
#define CF_METHOD_AUTO 2
CF_API int32_t cf_lowess_set_method(cf_lowess* h, int32_t method);
CF_API int32_t cf_lowess_method(const cf_lowess* h);
CF_API int32_t cf_set_default_method(int32_t method);
CF_API int32_t cf_default_method(void);

/* delta of later runs on this handle (default 0 = every distinct x is fitted).
 * delta > 0: single series (no batched handles).  On a range-sharded handle every rank
 * derives the whole fitted-point list (it must hold all of x: cf_lowess_init, not
 * cf_lowess_init_span) and fits the points of its block plus the fitted point before and
 * after it, so the interpolation needs no exchange.
 * Evaluator: with the direct kernel a tile is 512 consecutive FITTED points, whose windows lie
 * delta apart, so a tile sweeps far more elements than any of its points needs; for large delta
 * choose CF_METHOD_TREE / CF_METHOD_AUTO (n = 1e8 with 2 002 fitted points: 21 ms per smooth). */
CF_API int32_t cf_lowess_set_delta(cf_lowess* h, double delta);
CF_API double cf_lowess_delta(const cf_lowess* h);
/* After a run with delta > 0: number of points lowest() was called on and their
 * 1-based indices (first min(count, cap) of them; idx may be NULL). */
CF_API int32_t cf_lowess_fitted(cf_lowess* h, int32_t* count, int32_t* idx, int32_t cap);

/* The neighbourhoods (windows.cu) depend only on x and f, so a handle keeps them
 * between runs with the same f.  This drops them: the next run recomputes them,
 * like every call of the reference's lowess does (:658-669).  bench.py calls it
 * before each timed step so that a step is a COMPLETE smooth(). */
CF_API int32_t cf_lowess_reset_windows(cf_lowess* h);

/* ls_smooth's work on the device (lowess, :625-715), nsteps robustness
 * iterations, whole series on this GPU.  Synchronous.  stats may be NULL. */
CF_API int32_t cf_lowess_run(cf_lowess* h, double f, int32_t nsteps, cf_stats* stats);

/* Teacher forcing: `passes` fit passes starting from the given robustness
 * weights (host, length n; NULL = first pass without weights). */
CF_API int32_t cf_lowess_run_from_rw(cf_lowess* h, double f, const double* rw_in,
                                     int32_t passes);

/* ONE lowest() pass (:513-586 over the points of :662-673) of the DIRECT kernel on a
 * strided sample of its 512-point output tiles: tiles first_tile, first_tile +
 * tile_stride, ... (ntiles of them, clipped to the handle's shard), each evaluated
 * exactly as in a full pass (whole window, no window split -- the geometry a full
 * pass uses once the tiles alone fill the GPU).  use_rw = 0: the unweighted first
 * pass (:671 userw = .false.); 1: with the handle's current robustness weights.  ys
 * is written at the sampled points only; *ms (may be NULL) receives the CUDA-event
 * time of the launch.  This is how a direct pass at n = 1e8, ns = 5e6 -- 5.5 minutes
 * in full -- is measured and parity-checked in seconds.  Single series, delta = 0. */
CF_API int32_t cf_lowess_fit_sample(cf_lowess* h, double f, int32_t use_rw,
                                    int32_t first_tile, int32_t tile_stride,
                                    int32_t ntiles, float* ms);

/* Copy results to host buffers (each may be NULL): last ys, m_weights,
 * m_residuals (what get_residuals returns, :928-939). */
CF_API int32_t cf_lowess_fetch(cf_lowess* h, double* ys, double* rw, double* res);
/* Sorted x / y (what get_x / get_y index, :893-920). */
CF_API int32_t cf_lowess_fetch_xy(cf_lowess* h, double* x_sorted, double* y_sorted);
CF_API int32_t cf_lowess_count(const cf_lowess* h);  /* get_count, :876-884 */
/* Kernels of this library launched through the handle since it was created. */
CF_API int64_t cf_lowess_launch_count(const cf_lowess* h);

/* Neighbourhoods for span fraction f, 1-based, each array length n or NULL:
 * nleft/nright as the slide loop leaves them (:663-669); lo9/hi9 = first/last
 * index with a non-zero tricube weight (r <= .999h, :546); lo1/hi1 = the
 * w = 1 core (r <= .001h, :547-551); leader = index of the first member of the
 * point's tie group (:683-689; followers copy its ys). */
CF_API int32_t cf_lowess_windows(cf_lowess* h, double f,
                                 int32_t* nleft, int32_t* nright,
                                 int32_t* lo9, int32_t* hi9,
                                 int32_t* lo1, int32_t* hi1, int32_t* leader);

/* ------------------------------------------------------------------------
 * Stage API -- the same device stages cf_lowess_run chains, exposed so that a
 * multi-GPU driver (one process per GPU) can put its collectives between
 * them.  All stages are asynchronous on the handle's stream.
 * ---------------------------------------------------------------------- */
CF_API int32_t cf_lowess_stage_begin(cf_lowess* h, double f);        /* windows          */
CF_API int32_t cf_lowess_stage_fit(cf_lowess* h, int32_t use_rw);    /* fit + res, shard */
/* The select of the two middle order statistics of |res| (:697-703), one radix digit per
 * round, most significant first: round 0 counts the top 15 key bits of every residual of
 * the shard into CF_BUF_HIST0; rounds 1 .. cf_lowess_select_rounds() - 1 count 12 more
 * bits each, of the shard's CANDIDATES only (the keys of the bucket round 0 chose), into
 * CF_BUF_HIST.  The host runtime sums the counters over the ranks (integer all-reduce)
 * between _stage_hist and _stage_pick of the same round; the last pick forms cmad. */
CF_API int32_t cf_lowess_stage_hist(cf_lowess* h, int32_t round);    /* |res| digit hist */
CF_API int32_t cf_lowess_stage_pick(cf_lowess* h, int32_t round);    /* scan (post-sum)  */
CF_API int32_t cf_lowess_stage_weights(cf_lowess* h);                /* bisquare, shard  */
CF_API int32_t cf_lowess_select_rounds(void);
CF_API int32_t cf_lowess_sync(cf_lowess* h);

/* ------------------------------------------------------------------------
 * Peer-memory exchange (NVLink / NVSwitch; one process per GPU of one box).
 * Once attached, the robustness step of a range-sharded series needs no host
 * collective: cf_lowess_stage_pick sums the |res| digit histograms of all ranks
 * through peer stores inside the kernel that scans them (the global median,
 * :697-703), cf_lowess_stage_weights stores each new robustness weight into the
 * vector of every rank whose windows reach it (:704-713; the (ns-1)-point halo
 * exchange) and ends with a device-side barrier, cf_lowess_stage_gather
 * replicates the results.  Every call that runs passes on an attached handle
 * (the stage calls, cf_lowess_run) is collective: all ranks must make it.
 *   1. every rank: cf_lowess_init, cf_lowess_set_shard, cf_lowess_peer_export
 *      -> cf_lowess_peer_handle_bytes() bytes of CUDA IPC handles;
 *   2. the host runtime all-gathers them (rank-major) and every rank calls
 *      cf_lowess_peer_attach, then synchronises with the others once;
 *   3. all ranks issue the same sequence of stage calls.
 * A rank that never arrives does not hang the others: waits give up after
 * cf_lowess_peer_set_timeout seconds (default 60), cf_lowess_run and
 * cf_lowess_stage_gather then return CF_DEVICE_ERROR (cf_lowess_peer_error reads the same
 * flag).  cf_lowess_set_shard with a different layout detaches; cf_lowess_init on an
 * attached handle keeps the mapping when the buffers stay in place (same n) and FAILS with
 * CF_INVALID_INPUT_ERROR when it would have to reallocate them -- detach on all ranks
 * first (a peer may still be storing into the old buffers).  At most 8 ranks.
 * ---------------------------------------------------------------------- */
CF_API int32_t cf_lowess_peer_handle_bytes(void);
CF_API int32_t cf_lowess_peer_export(cf_lowess* h, void* handles_out);
CF_API int32_t cf_lowess_peer_attach(cf_lowess* h, const void* all_handles, int32_t nranks);
/* Same-process variant: hs[r] holds shard r of nranks (one handle per device, all created
 * by this process); peers are reached through cudaDeviceEnablePeerAccess.  This is what
 * cf_lowess_smooth_multi uses. */
CF_API int32_t cf_lowess_peer_attach_local(cf_lowess* const* hs, int32_t nranks);
/* Device-side waits give up after `seconds` (default 60) and the run returns
 * CF_DEVICE_ERROR.  Set before attaching or at any time after. */
CF_API int32_t cf_lowess_peer_set_timeout(cf_lowess* h, double seconds);
CF_API int32_t cf_lowess_peer_detach(cf_lowess* h);
CF_API int32_t cf_lowess_peer_attached(const cf_lowess* h);
CF_API int32_t cf_lowess_peer_error(cf_lowess* h, int32_t* flag);
/* Replicate the own block of ys (CF_GATHER_YS), res (CF_GATHER_RES), rw (CF_GATHER_RW;
 * during the run a rank receives only the weights its windows reach) on all ranks. */
#define CF_GATHER_YS 1
#define CF_GATHER_RES 2
#define CF_GATHER_RW 4
CF_API int32_t cf_lowess_stage_gather(cf_lowess* h, int32_t what);

/* Device pointers of the handle's buffers, for zero-copy wrapping by the host
 * runtime (e.g. as torch tensors for NCCL collectives). */
#define CF_BUF_X 0
#define CF_BUF_Y 1
#define CF_BUF_YS 2
#define CF_BUF_RW 3      /* length = nranks * block (padded for all-gather)  */
#define CF_BUF_RES 4
#define CF_BUF_HIST 5    /* uint32 counters, cf_lowess_hist_len() of them     */
#define CF_BUF_CMAD 6    /* double[3]: cmad, c1, c9 of the last iteration     */
#define CF_BUF_HIST0 7   /* uint32 counters, cf_lowess_hist0_len() of them    */
CF_API void* cf_lowess_buffer(cf_lowess* h, int32_t which);
CF_API int32_t cf_lowess_hist_len(void);
CF_API int32_t cf_lowess_hist0_len(void);

/* FP64 pipe peak microbenchmark (independent DFMA chains on every SM):
 * returns measured TFLOP/s (FMA = 2 flop) over `ms_target` milliseconds. */
CF_API int32_t cf_fp64_peak(int32_t device, double ms_target, double* tflops,
                            double* sm_mhz_est);

#ifdef __cplusplus
}
#endif
#endif /* CURVEFIT_CUDA_H_ */
