// Shared device helpers of the in-kernel NVLink collectives (symm.cu, symm_fused.cu): peer pointer table, system-scope
// release/acquire flag accesses, epoch signal/wait. See symm.cu for the protocol.
#pragma once
#include "common.cuh"

namespace {

constexpr int MAX_WORLD = 16;
constexpr int FLAG_STRIDE = 32;  // uint32 slots per channel (>= MAX_WORLD)

struct PeerPtrs {
    void* p[MAX_WORLD];
};

__device__ __forceinline__ void st_release_sys(uint32_t* addr, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* addr)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p)
{
    uint4 r;
    asm volatile("ld.global.relaxed.sys.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p)
                 : "memory");
    return r;
}

// Tell every peer (including myself) that I reached `epoch` on `channel`.
__device__ __forceinline__ void signal_all(const PeerPtrs& flags, int rank, int world, int channel, uint32_t epoch)
{
    if (threadIdx.x < world) {
        uint32_t* f = reinterpret_cast<uint32_t*>(flags.p[threadIdx.x]) + channel * FLAG_STRIDE + rank;
        st_release_sys(f, epoch);
    }
}
// Wait until every peer has reached `epoch` on `channel` (reads my local flag words only).
__device__ __forceinline__ void wait_all(const PeerPtrs& flags, int rank, int world, int channel, uint32_t epoch)
{
    if (threadIdx.x < world) {
        const uint32_t* f = reinterpret_cast<const uint32_t*>(flags.p[rank]) + channel * FLAG_STRIDE + threadIdx.x;
        while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) { __nanosleep(64); }
    }
    __syncthreads();
}

inline PeerPtrs make_peer_ptrs(const uint64_t* p, int world)
{
    PeerPtrs r;
    for (int i = 0; i < MAX_WORLD; i++) r.p[i] = i < world ? reinterpret_cast<void*>(p[i]) : nullptr;
    return r;
}

}  // namespace
