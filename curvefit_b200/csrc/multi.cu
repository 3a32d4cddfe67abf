// multi.cu -- cf_lowess_smooth_multi: lowess() (src/curvefit_regression.f90:625-715, as
// ls_smooth calls it, :866) of ONE series over several GPUs of one box from ONE process --
// what a Fortran program calling fit%smooth(f) can reach through the bind(C) shim.
//
// The sorted index range is cut into `ndev` equal blocks of output points (same
// partition and the same kernels as the one-process-per-GPU path of distributed.py).  One
// host thread per device:
//   * uploads only the elements its block's windows can reach -- nleft(begin) .. nright(end - 1) + 1
//     widened over ties: at most the (ns-1)-point halo either side -- not the whole series;
//   * runs the passes (cf_lowess_run on a peer-attached handle: the histogram all-reduce of the
//     median, the candidate exchange and the halo weights cross NVLink inside the kernels,
//     peer.cu; the peers are mapped with cudaDeviceEnablePeerAccess, no IPC);
//   * copies its own block of ys, rw, res straight into the caller's arrays.
// The handles, their buffers and the peer mapping are kept between calls of the same n and
// device list (a caller looping over smooth() pays for them once); cf_lowess_multi_release()
// frees them.
#include <algorithm>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "common.cuh"

namespace cf {
namespace {

struct MultiCtx {
    std::vector<int> devs;
    int32_t n = 0;
    std::vector<cf_lowess*> hs;
};

std::mutex g_multi_mu;            // one multi-GPU smooth at a time per process
MultiCtx* g_multi = nullptr;
std::vector<int> g_no_peer_devs;  // a device list whose peer mapping failed: peer access does not change
                                  // while the process lives, so later calls go to one GPU at once

void multi_free(MultiCtx* c)
{
    if (!c) return;
    for (cf_lowess* h : c->hs) if (h) { cudaSetDevice(h->device); cudaStreamSynchronize(h->stream); }
    for (cf_lowess* h : c->hs) if (h) { cudaSetDevice(h->device); peer_detach(h); }
    for (cf_lowess* h : c->hs) cf_lowess_destroy(h);
    delete c;
}

// *no_peers: the per-device buffers fitted but the devices cannot map each other's memory
int multi_build(const std::vector<int>& devs, int32_t n, MultiCtx** out, bool* no_peers)
{
    *no_peers = false;
    MultiCtx* c = new MultiCtx();
    c->devs = devs;
    c->n = n;
    c->hs.assign(devs.size(), nullptr);
    int rc = CF_OK;
    for (size_t r = 0; r < devs.size() && rc == CF_OK; ++r) {
        rc = cf_lowess_create(&c->hs[r], devs[r], nullptr);
        if (rc == CF_OK) rc = cf_lowess_set_shard(c->hs[r], (int32_t)r, (int32_t)devs.size());
        if (rc == CF_OK) rc = layout_series(c->hs[r], n);
    }
    if (rc == CF_OK) {
        rc = cf_lowess_peer_attach_local(c->hs.data(), (int32_t)devs.size());
        *no_peers = rc != CF_OK;
    }
    if (rc != CF_OK) {
        multi_free(c);
        for (int d : devs) { if (cudaSetDevice(d) == cudaSuccess) (void)cudaGetLastError(); }   // a failed cudaMalloc is not sticky
        return rc;
    }
    *out = c;
    return CF_OK;
}

// all threads of one call meet here between the upload and the collective run
struct HostBarrier {
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0, total = 0, gen = 0;
    void arrive()
    {
        std::unique_lock<std::mutex> lk(mu);
        const int g = gen;
        if (++waiting == total) { waiting = 0; ++gen; cv.notify_all(); }
        else cv.wait(lk, [&] { return gen != g; });
    }
};

}  // namespace
}  // namespace cf

using namespace cf;

extern "C" {

int32_t cf_lowess_multi_release(void)
{
    std::lock_guard<std::mutex> lk(g_multi_mu);
    int dev = 0;
    const bool have = cudaGetDevice(&dev) == cudaSuccess;
    multi_free(g_multi);
    g_multi = nullptr;
    if (have) cudaSetDevice(dev);
    return CF_OK;
}

int32_t cf_lowess_multi_active(void)
{
    std::lock_guard<std::mutex> lk(g_multi_mu);
    return g_multi ? (int32_t)g_multi->devs.size() : 0;
}

int32_t cf_lowess_smooth_multi(const double* x_sorted, const double* y, int32_t n, double f,
                               int32_t nsteps, const int32_t* devices, int32_t ndev,
                               double* ys, double* rw, double* res)
{
    if (n < 0 || (n > 0 && (!x_sorted || !y || !ys))) {
        set_error("Input array sizes must match.");
        return CF_ARRAY_SIZE_ERROR;
    }
    if (nsteps < 0) { set_error("nsteps must be >= 0"); return CF_INVALID_INPUT_ERROR; }
    int visible = 0;
    {
        const cudaError_t e = cudaGetDeviceCount(&visible);
        if (e != cudaSuccess || visible == 0) {
            set_error("no CUDA device available (%s); libcurvefit_cuda has no CPU fallback",
                      e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
            return CF_DEVICE_ERROR;
        }
    }
    std::vector<int> devs;
    if (!devices || ndev <= 0) {
        for (int d = 0; d < visible && d < kMaxPeers; ++d) devs.push_back(d);
    } else {
        if (ndev > kMaxPeers) { set_error("at most %d devices", kMaxPeers); return CF_INVALID_INPUT_ERROR; }
        for (int k = 0; k < ndev; ++k) {
            if (devices[k] < 0 || devices[k] >= visible || std::count(devs.begin(), devs.end(), (int)devices[k])) {
                set_error("bad device list (entry %d = %d; %d visible, no duplicates)", k, (int)devices[k], visible);
                return CF_INVALID_INPUT_ERROR;
            }
            devs.push_back(devices[k]);
        }
    }
    int caller_dev = 0;
    CF_CUDA(cudaGetDevice(&caller_dev));
    // short series: the blocks would be a few tiles each -- one GPU does it
    if (devs.size() == 1 || (int64_t)n < (int64_t)4096 * (int64_t)devs.size()) {
        CF_CUDA(cudaSetDevice(devs[0]));
        const int rc = cf_lowess_smooth(x_sorted, y, n, f, nsteps, ys, rw, res);
        cudaSetDevice(caller_dev);
        return rc;
    }
    if (!(f == f)) { set_error("f is NaN"); return CF_INVALID_INPUT_ERROR; }

    // The sharded run needs room for the buffers on every device and a peer mapping between all of
    // them.  A box without P2P (PCIe without a switch, mixed GPUs, MIG slices) or a GPU another job
    // has filled must not turn a smooth() that works on one GPU into an error: the context is
    // freed and the whole series goes to ONE device -- the caller's if it is in the list, else the
    // first of the list -- with identical results (cf_lowess_smooth).
    auto one_gpu = [&]() -> int32_t {
        const int d = std::count(devs.begin(), devs.end(), caller_dev) ? caller_dev : devs[0];
        CF_CUDA(cudaSetDevice(d));
        const int rc = cf_lowess_smooth(x_sorted, y, n, f, nsteps, ys, rw, res);
        cudaSetDevice(caller_dev);
        return rc;
    };
    std::unique_lock<std::mutex> lk(g_multi_mu);
    if (g_multi && (g_multi->n != n || g_multi->devs != devs)) { multi_free(g_multi); g_multi = nullptr; }
    if (!g_multi && g_no_peer_devs == devs) { lk.unlock(); return one_gpu(); }
    if (!g_multi) {
        bool no_peers = false;
        const int rc = multi_build(devs, n, &g_multi, &no_peers);
        if (rc != CF_OK) {
            g_multi = nullptr;
            if (no_peers) g_no_peer_devs = devs;
            lk.unlock();
            return one_gpu();
        }
    }
    MultiCtx* c = g_multi;
    const int N = (int)devs.size();
    const int32_t ns = cf_lowess_span(n, f);
    const int method = stateless_method();      // read on the caller's thread, handed to the rank threads
    std::vector<int> rcs((size_t)N, CF_OK);
    std::vector<std::string> msgs((size_t)N);
    HostBarrier bar;
    bar.total = N;
    bool abort_run = false;       // written by rank threads before the barrier, read after it

    auto work = [&](int r) {
        cf_lowess* h = c->hs[(size_t)r];
        auto fail = [&](int rc) { rcs[(size_t)r] = rc; msgs[(size_t)r] = last_error_text(); };
        int rc = CF_OK;
        if (cudaSetDevice(h->device) != cudaSuccess) rc = CF_DEVICE_ERROR;
        h->method = method;
        // only the elements the windows of [begin, end) can reach (shard_span_host, api.cu)
        int32_t lo = 0, hi = 0;
        shard_span_host(x_sorted, n, ns, h->begin, h->end, &lo, &hi);
#ifdef CF_DEVELOPMENT
        const auto t0 = std::chrono::steady_clock::now();
        auto since = [&]() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); };
#endif
        if (rc == CF_OK) rc = upload_span(h, x_sorted, y, lo, hi);
#ifdef CF_DEVELOPMENT
        const double t_enq = since();
        cudaStreamSynchronize(h->stream);
        cudaStreamSynchronize(h->copy_stream);
        const double t_up = since();
#endif
        // Everything the run allocates (candidate list, partial moments / tree nodes) is reserved
        // here, before the barrier: a rank whose cudaMalloc fails must say so while the others can
        // still stay out of the collective run -- once they are inside, they would spin in the
        // device-side peer waits until the timeout, holding g_multi_mu.
        if (rc == CF_OK) rc = select_prepare(h);
        if (rc == CF_OK) {
            h->ns = ns;
            h->reserve_only = true;
            rc = launch_fit(h, false);
            h->reserve_only = false;
        }
        if (rc != CF_OK) {
            fail(rc);
            std::lock_guard<std::mutex> g(bar.mu);
            abort_run = true;
        }
        bar.arrive();             // nobody enters the collective run unless every rank can
        if (abort_run) return;
        h->early_rw = (rw && nsteps > 0) ? rw : nullptr;      // leaves beside the last fit pass
        rc = cf_lowess_run(h, f, nsteps, nullptr);
        if (rc == CF_OK) {
            const size_t cnt = (size_t)(h->end - h->begin), b = sizeof(double) * cnt;
            cudaError_t e = cudaMemcpyAsync(ys + h->begin, h->ys + h->begin, b, cudaMemcpyDeviceToHost, h->stream);
            if (e == cudaSuccess && h->early_rw_missed)
                e = cudaMemcpyAsync(rw + h->begin, h->rw + h->begin, b, cudaMemcpyDeviceToHost, h->stream);
            h->early_rw_missed = nullptr;
            if (e == cudaSuccess && res)
                e = cudaMemcpyAsync(res + h->begin, h->res + h->begin, b, cudaMemcpyDeviceToHost, h->stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
            if (e == cudaSuccess && h->copy_stream) e = cudaStreamSynchronize(h->copy_stream);
            if (e != cudaSuccess) rc = cuda_fail(e, "cf_lowess_smooth_multi download", __FILE__, __LINE__);
        }
        if (rc != CF_OK) fail(rc);
#ifdef CF_DEVELOPMENT
        if (dev_env("CF_MULTI_TIMES"))
            fprintf(stderr, "[multi] rank %d dev %d: upload enqueued %.2f ms, landed %.2f ms ([%d, %d] = %.0f MB), all done %.2f ms\n",
                    r, h->device, t_enq, t_up, lo, hi, 16e-6 * (hi - lo + 1), since());
#endif
    };
    std::vector<std::thread> th;
    for (int r = 1; r < N; ++r) th.emplace_back(work, r);
    work(0);
    for (auto& t : th) t.join();
    cudaSetDevice(caller_dev);
    if (abort_run) {
        // nobody entered the collective run.  If all that went wrong is memory (the evaluator's
        // scratch did not fit beside the buffers on some device), one GPU gets the whole series,
        // like when the context itself cannot be built.
        bool only_memory = true;
        for (int r = 0; r < N; ++r)
            if (rcs[(size_t)r] != CF_OK && rcs[(size_t)r] != CF_OUT_OF_MEMORY_ERROR) only_memory = false;
        if (only_memory) {
            multi_free(g_multi);
            g_multi = nullptr;
            for (int d : devs) { if (cudaSetDevice(d) == cudaSuccess) (void)cudaGetLastError(); }
            lk.unlock();
            return one_gpu();
        }
    }
    for (int r = 0; r < N; ++r) {
        if (rcs[(size_t)r] != CF_OK) {
            set_error("device %d (block %d of %d): %s", devs[(size_t)r], r, N, msgs[(size_t)r].c_str());
            // a failed collective leaves epochs and flags in an unknown state: start afresh next time
            multi_free(g_multi);
            g_multi = nullptr;
            cudaSetDevice(caller_dev);
            return rcs[(size_t)r];
        }
    }
    return CF_OK;
}

}  // extern "C"
