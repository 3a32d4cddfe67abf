// peer.cuh -- device primitives of the peer-memory exchange (see peer.cu).
#pragma once

#include "common.cuh"

namespace cf {

constexpr unsigned long long kPeerTimeoutDefaultNs = 60000000000ull;   // 60 s, then PeerComm::error

__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned int ld_relaxed_sys_u32(const unsigned int* p)
{
    unsigned int v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Called by the first kMaxPeers threads of a CTA after all of the CTA's peer stores
// are fenced (caller: __threadfence_system(); __syncthreads()).  Thread p raises this
// rank's epoch in rank p's PeerComm and waits for rank p's epoch in its own.
__device__ __forceinline__ void peer_signal_wait(const PeerPtrs& pp, int channel,
                                                 unsigned long long epoch)
{
    const int p = threadIdx.x;
    if (p < pp.nranks) {
        st_release_sys(&pp.comm[p]->flag[channel][pp.rank], epoch);
        const unsigned long long* mine = &pp.comm[pp.rank]->flag[channel][p];
        const unsigned long long t0 = global_ns();
        while (ld_acquire_sys(mine) < epoch) {
            if (global_ns() - t0 > pp.timeout_ns) {
                pp.comm[pp.rank]->error = 1;
                break;
            }
            __nanosleep(64);
        }
    }
}

// All-reduce (sum) of the `len` <= 2^kSelB0Single counters hist[] over the ranks, result in place
// on every rank.  Called by ALL threads of a single CTA.
__device__ __forceinline__ void peer_hist_allreduce(const PeerPtrs& pp, unsigned int* hist, int len,
                                                    unsigned long long epoch)
{
    const int par = (int)(epoch & 1ull);
    for (int k = threadIdx.x; k < len; k += blockDim.x) {
        const unsigned int v = __ldcg(hist + k);
#pragma unroll
        for (int p = 0; p < kMaxPeers; ++p)
            if (p < pp.nranks) pp.comm[p]->slots[par][pp.rank][k] = v;
    }
    __threadfence_system();
    __syncthreads();
    peer_signal_wait(pp, 0, epoch);
    __syncthreads();
    const PeerComm* own = pp.comm[pp.rank];
    for (int k = threadIdx.x; k < len; k += blockDim.x) {
        unsigned int s = 0u;
        for (int p = 0; p < pp.nranks; ++p) s += ld_relaxed_sys_u32(&own->slots[par][p][k]);
        hist[k] = s;
    }
    __threadfence();
    __syncthreads();
}

}  // namespace cf
