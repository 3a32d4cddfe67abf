// peer.cu -- the two exchanges of a range-sharded robustness step done
// by the kernels themselves over NVLink peer memory, not by NCCL:
//
//   * global median of |res| (src/curvefit_regression.f90:697-703): the last CTA of the
//     fused residual + round-0 histogram pass (robust.cu) stores this rank's 2^15 digit
//     counters into slot [rank] of EVERY rank's PeerComm, raises its epoch flag there,
//     waits for the flags of all ranks in its own PeerComm and sums the slots -- an integer
//     all-reduce fused into the scan that consumes it; the keys of the bucket that holds the
//     median (1-3 % of the shard) are then stored into every rank's candidate list
//     (peer_push_kernel), so the remaining rounds of the select run on every rank
//     without another exchange and end in the same cmad, exactly;
//   * refreshed robustness weights (:704-713): bisquare_kernel<true> stores each
//     weight of its block into the rw vector of every rank whose windows can reach
//     it -- the (ns-1)-point HALO exchange: each rank publishes the element span
//     [min lo9, max hi9] of its shard's windows once per smooth (peer_spans_kernel) --
//     followed by peer_barrier_kernel so that no rank starts the next fit pass before
//     its halo has landed;
//   * peer_gather_kernel stores the own block of ys (and rw, when asked for) into every
//     rank at the end; res of the other blocks is completed locally (every rank holds y).
//
// Memory is shared between the one-process-per-GPU ranks with CUDA IPC handles
// (cf_lowess_peer_export / _attach; the host runtime moves the 64-byte handles, e.g.
// with torch.distributed.all_gather).  Flags are monotone epochs, so a slot is reused
// only two rounds later, after every rank has passed the round in between.  All waits
// give up after kPeerTimeoutNs and set PeerComm::error instead of hanging the GPU.
#include <string.h>

#include "common.cuh"
#include "peer.cuh"

namespace cf {

__global__ void __launch_bounds__(32)
peer_barrier_kernel(const PeerPtrs pp, unsigned long long epoch)
{
    peer_signal_wait(pp, 1, epoch);
}

// Publishes this rank's element span, meets the other ranks, copies their spans.
__global__ void __launch_bounds__(32)
peer_spans_kernel(const PeerPtrs pp, const int32_t* __restrict__ span, unsigned long long epoch)
{
    PeerComm* own = pp.comm[pp.rank];
    if (threadIdx.x == 0) {
        own->span[0] = span[0];
        own->span[1] = span[1];
    }
    __threadfence_system();
    __syncwarp();
    peer_signal_wait(pp, 1, epoch);
    const int p = threadIdx.x;
    if (p < pp.nranks) {
        const volatile int32_t* theirs = pp.comm[p]->span;
        own->peer_span[p][0] = theirs[0];
        own->peer_span[p][1] = theirs[1];
    }
}

// Own candidates (slot `rank` of the local list, st->ncand of them) -> the same slot of every
// peer's list; the last CTA to finish publishes the count everywhere and meets the other
// ranks, after which every rank holds the candidates of all ranks.
__global__ void __launch_bounds__(256)
peer_push_kernel(const PeerPtrs pp, SelectState* st, int64_t slot_cap, unsigned long long epoch)
{
    __shared__ unsigned int sh_last;
    const unsigned int cnt = st->ncand;
    const unsigned long long* __restrict__ mine = pp.cand[pp.rank] + (size_t)pp.rank * (size_t)slot_cap;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < cnt; k += stride) {
        const unsigned long long v = mine[k];
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
            if (p < pp.nranks && p != pp.rank) pp.cand[p][(size_t)pp.rank * (size_t)slot_cap + k] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) sh_last = (atomicAdd(&st->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!sh_last) return;
    __threadfence_system();
    if (threadIdx.x < pp.nranks) pp.comm[threadIdx.x]->cand_count[pp.rank] = cnt;
    if (threadIdx.x == 0) st->ticket = 0u;
    __threadfence_system();
    __syncthreads();
    peer_signal_wait(pp, 1, epoch);
}

// own block [begin, end) of ys (and, if wanted, of rw -- the weights of the LAST robustness
// iteration, which are not a function of the final residuals) -> the same positions on
// every other rank.  res does not cross NVLink: every rank holds all of y and completes
// res = y - ys (:695) for the other blocks itself (peer_complete_kernel), bit-identically.
__global__ void __launch_bounds__(256)
peer_gather_kernel(const PeerPtrs pp, int32_t begin, int32_t end, int what)
{
    const int32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= end) return;
    const double v = pp.ys[pp.rank][i], w = pp.rw[pp.rank][i];
#pragma unroll
    for (int p = 0; p < kMaxPeers; ++p) {
        if (p < pp.nranks && p != pp.rank) {
            pp.ys[p][i] = v;
            if (what & CF_GATHER_RW) pp.rw[p][i] = w;
        }
    }
}

__global__ void __launch_bounds__(256)
peer_complete_kernel(const double* __restrict__ y, const double* __restrict__ ys, double* __restrict__ res,
                     int32_t begin, int32_t end, int32_t n)
{
    int32_t i = blockIdx.x * blockDim.x + threadIdx.x;       // index among the points NOT in [begin, end)
    if (i >= begin) i += end - begin;
    if (i >= n) return;
    res[i] = __dsub_rn(y[i], ys[i]);
}

int launch_peer_barrier(cf_lowess* h)
{
    if (!h->peer_attached) return CF_OK;
    peer_barrier_kernel<<<1, 32, 0, h->stream>>>(h->peer, ++h->epoch_bar);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_peer_spans(cf_lowess* h)
{
    if (!h->peer_attached) return CF_OK;
    peer_spans_kernel<<<1, 32, 0, h->stream>>>(h->peer, h->span, ++h->epoch_bar);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_peer_push(cf_lowess* h)
{
    if (!h->peer_attached) return CF_OK;
    const int64_t cnt = h->end > h->begin ? h->end - h->begin : 0;
    int64_t blocks = cnt / 32 / 256 / 4 + 1;        // ~3 % of the shard are candidates
    if (blocks > 148) blocks = 148;
    peer_push_kernel<<<(int)blocks, 256, 0, h->stream>>>(h->peer, h->sel, (int64_t)h->block, ++h->epoch_bar);
    CF_CUDA(cudaGetLastError());
    h->launches++;
    return CF_OK;
}

int launch_peer_gather(cf_lowess* h, int what)
{
    if (!h->peer_attached) { set_error("peers are not attached"); return CF_INVALID_INPUT_ERROR; }
    const int cnt = h->end - h->begin;
    if (cnt > 0 && (what & 7)) {
        peer_gather_kernel<<<(cnt + 255) / 256, 256, 0, h->stream>>>(h->peer, h->begin, h->end, what);
        CF_CUDA(cudaGetLastError());
        h->launches++;
    }
    CF_TRY(launch_peer_barrier(h));
    const int rest = h->n - cnt;
    if (rest > 0 && (what & CF_GATHER_RES)) {
        peer_complete_kernel<<<(rest + 255) / 256, 256, 0, h->stream>>>(h->y, h->ys, h->res, h->begin, h->end,
                                                                        h->n);
        CF_CUDA(cudaGetLastError());
        h->launches++;
    }
    return CF_OK;
}

// Buffers the peers map: the exchange block (zeroed, epochs restarted) and the candidate list.
int peer_prepare(cf_lowess* h)
{
    peer_detach(h);
    CF_TRY(select_prepare(h));
    if (!h->comm) CF_CUDA(cudaMalloc(&h->comm, sizeof(PeerComm)));
    CF_CUDA(cudaMemsetAsync(h->comm, 0, sizeof(PeerComm), h->stream));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    h->epoch_hist = h->epoch_bar = 0;
    return CF_OK;
}

// After a run on an attached handle: a wait that timed out means ranks worked with missing
// histograms / weights -- the results are not to be trusted.
int peer_check(cf_lowess* h)
{
    if (!h->peer_attached || !h->comm) return CF_OK;
    int v = 0;
    CF_CUDA(cudaMemcpyAsync(&v, &h->comm->error, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    if (v) {
        set_error("a peer wait timed out (a rank did not reach the exchange within %.1f s)", (double)h->peer_timeout_ns * 1e-9);
        return CF_DEVICE_ERROR;
    }
    return CF_OK;
}

int peer_clear_error(cf_lowess* h)
{
    if (!h->peer_attached || !h->comm) return CF_OK;
    CF_CUDA(cudaMemsetAsync(&h->comm->error, 0, sizeof(int), h->stream));
    return CF_OK;
}

void peer_detach(cf_lowess* h)
{
    if (!h->peer_attached) return;
    cudaStreamSynchronize(h->stream);
    for (int p = 0; p < h->peer.nranks && h->peer_ipc; ++p) {
        if (p == h->peer.rank) continue;
        if (h->peer.comm[p]) cudaIpcCloseMemHandle(h->peer.comm[p]);
        if (h->peer.rw[p]) cudaIpcCloseMemHandle(h->peer.rw[p]);
        if (h->peer.ys[p]) cudaIpcCloseMemHandle(h->peer.ys[p]);
        if (h->peer.res[p]) cudaIpcCloseMemHandle(h->peer.res[p]);
        if (h->peer.cand[p]) cudaIpcCloseMemHandle(h->peer.cand[p]);
    }
    h->peer = PeerPtrs{};
    h->peer_attached = false;
}

}  // namespace cf

using namespace cf;

extern "C" {

int32_t cf_lowess_peer_handle_bytes(void) { return 5 * (int32_t)sizeof(cudaIpcMemHandle_t); }

int32_t cf_lowess_peer_export(cf_lowess* h, void* out)
{
    if (!h || !out) { set_error("null argument"); return CF_INVALID_INPUT_ERROR; }
    if (!h->init || h->nseries != 1) { set_error("No data has been defined."); return CF_NO_DATA_DEFINED_ERROR; }
    if (h->nranks > kMaxPeers) { set_error("at most %d ranks", kMaxPeers); return CF_INVALID_INPUT_ERROR; }
    CF_CUDA(cudaSetDevice(h->device));
    CF_TRY(peer_prepare(h));
    cudaIpcMemHandle_t hs[5];
    CF_CUDA(cudaIpcGetMemHandle(&hs[0], h->comm));
    CF_CUDA(cudaIpcGetMemHandle(&hs[1], h->rw));
    CF_CUDA(cudaIpcGetMemHandle(&hs[2], h->ys));
    CF_CUDA(cudaIpcGetMemHandle(&hs[3], h->res));
    CF_CUDA(cudaIpcGetMemHandle(&hs[4], h->cand));
    memcpy(out, hs, sizeof(hs));
    return CF_OK;
}

int32_t cf_lowess_peer_attach(cf_lowess* h, const void* all_handles, int32_t nranks)
{
    if (!h || !all_handles) { set_error("null argument"); return CF_INVALID_INPUT_ERROR; }
    if (nranks != h->nranks || nranks > kMaxPeers || !h->comm) {
        set_error("peer_attach: %d ranks given, shard has %d (export first; at most %d)", (int)nranks,
                  (int)h->nranks, kMaxPeers);
        return CF_INVALID_INPUT_ERROR;
    }
    CF_CUDA(cudaSetDevice(h->device));
    PeerPtrs pp = {};
    pp.rank = h->rank;
    pp.nranks = nranks;
    pp.timeout_ns = h->peer_timeout_ns;
    const cudaIpcMemHandle_t* hs = static_cast<const cudaIpcMemHandle_t*>(all_handles);
    int rc = CF_OK;
    for (int p = 0; p < nranks && rc == CF_OK; ++p) {
        if (p == h->rank) {
            pp.comm[p] = h->comm; pp.rw[p] = h->rw; pp.ys[p] = h->ys; pp.res[p] = h->res; pp.cand[p] = h->cand;
            continue;
        }
        void* ptr[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
        for (int k = 0; k < 5; ++k) {
            cudaError_t e = cudaIpcOpenMemHandle(&ptr[k], hs[5 * p + k], cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) { rc = cuda_fail(e, "cudaIpcOpenMemHandle", __FILE__, __LINE__); break; }
        }
        pp.comm[p] = static_cast<PeerComm*>(ptr[0]);
        pp.rw[p] = static_cast<double*>(ptr[1]);
        pp.ys[p] = static_cast<double*>(ptr[2]);
        pp.res[p] = static_cast<double*>(ptr[3]);
        pp.cand[p] = static_cast<unsigned long long*>(ptr[4]);
    }
    h->peer = pp;
    h->peer_attached = true;       // so that peer_detach closes what was opened
    h->peer_ipc = true;
    if (rc != CF_OK) peer_detach(h);
    return rc;
}

// Same-process variant: the peers are other handles of THIS process (one per device); their
// buffers are reached through cudaDeviceEnablePeerAccess instead of CUDA IPC.
int32_t cf_lowess_peer_attach_local(cf_lowess* const* hs, int32_t nranks)
{
    if (!hs || nranks < 1 || nranks > kMaxPeers) { set_error("peer_attach_local: 1..%d handles", kMaxPeers); return CF_INVALID_INPUT_ERROR; }
    for (int p = 0; p < nranks; ++p) {
        cf_lowess* h = hs[p];
        if (!h || !h->init || h->nseries != 1 || h->nranks != nranks || h->rank != p) {
            set_error("peer_attach_local: handle %d must be initialised and hold shard %d of %d", p, p, (int)nranks);
            return CF_INVALID_INPUT_ERROR;
        }
        CF_CUDA(cudaSetDevice(h->device));
        CF_TRY(peer_prepare(h));
    }
    for (int p = 0; p < nranks; ++p) {
        CF_CUDA(cudaSetDevice(hs[p]->device));
        for (int q = 0; q < nranks; ++q) {
            if (hs[q]->device == hs[p]->device) continue;
            int can = 0;
            CF_CUDA(cudaDeviceCanAccessPeer(&can, hs[p]->device, hs[q]->device));
            if (!can) { set_error("device %d cannot access device %d", hs[p]->device, hs[q]->device); return CF_DEVICE_ERROR; }
            const cudaError_t e = cudaDeviceEnablePeerAccess(hs[q]->device, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
        }
    }
    for (int p = 0; p < nranks; ++p) {
        cf_lowess* h = hs[p];
        PeerPtrs pp = {};
        pp.rank = p;
        pp.nranks = nranks;
        pp.timeout_ns = h->peer_timeout_ns;
        for (int q = 0; q < nranks; ++q) {
            pp.comm[q] = hs[q]->comm; pp.rw[q] = hs[q]->rw; pp.ys[q] = hs[q]->ys; pp.res[q] = hs[q]->res;
            pp.cand[q] = hs[q]->cand;
        }
        h->peer = pp;
        h->peer_attached = true;
        h->peer_ipc = false;
    }
    return CF_OK;
}

int32_t cf_lowess_peer_set_timeout(cf_lowess* h, double seconds)
{
    if (!h || !(seconds > 0.0) || !(seconds < 1.0e6)) { set_error("timeout must be in (0, 1e6) seconds"); return CF_INVALID_INPUT_ERROR; }
    h->peer_timeout_ns = (unsigned long long)(seconds * 1.0e9);
    h->peer.timeout_ns = h->peer_timeout_ns;
    return CF_OK;
}

int32_t cf_lowess_peer_detach(cf_lowess* h)
{
    if (!h) return CF_INVALID_INPUT_ERROR;
    cudaSetDevice(h->device);
    peer_detach(h);
    return CF_OK;
}

int32_t cf_lowess_peer_attached(const cf_lowess* h) { return h && h->peer_attached ? 1 : 0; }

int32_t cf_lowess_stage_gather(cf_lowess* h, int32_t what)
{
    if (!h || !h->init) return CF_NO_DATA_DEFINED_ERROR;
    CF_CUDA(cudaSetDevice(h->device));
    CF_TRY(launch_peer_gather(h, what));
    return peer_check(h);
}

int32_t cf_lowess_peer_error(cf_lowess* h, int32_t* flag)
{
    if (!h || !flag) return CF_INVALID_INPUT_ERROR;
    *flag = 0;
    if (!h->comm) return CF_OK;
    CF_CUDA(cudaSetDevice(h->device));
    int v = 0;
    CF_CUDA(cudaMemcpyAsync(&v, &h->comm->error, sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    CF_CUDA(cudaStreamSynchronize(h->stream));
    *flag = v;
    if (v) set_error("a peer wait timed out (a rank did not reach the exchange)");
    return CF_OK;
}

}  // extern "C"
