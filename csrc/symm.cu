// In-kernel NVLink 5 / NVSwitch collectives over symmetric (peer-mapped) memory.
//
// Reference baseline: every collective of the reference is an NCCL call between kernels
// (coalesced_collectives.py:158-218 reduce_scatter_coalesced; partition_parameters.py:1301-1343 all-gather;
// sequence/layer.py:264 all-to-all; sharded_moe.py:104 all-to-all). The closest it gets to peer memory is a
// barrier + per-peer copy_ loop (csrc/compile/z3.cpp:95-113).
// Here the data movement is *inside* the kernel: each CTA issues 16-byte loads from the peers' buffers through
// their NVLink-mapped addresses (or one multimem.ld_reduce / multimem.st through the NVSwitch multicast address)
// and fuses the adjacent elementwise work:
//   * reduce-scatter  + average/unscale + cast + accumulate-into-shard + sum-of-squares partial   (ZeRO path 1)
//   * all-gather      (pull from peers, or multimem.st push)                                       (ZeRO path 2)
//   * all-to-all      with arbitrary chunk layout (Ulysses head/sequence re-shard, MoE dispatch)   (paths 4/5)
// Cross-GPU ordering uses monotonically increasing epoch flags in a symmetric flag buffer:
//   flags[FLAG_READY + src]  : "src has finished writing epoch e"   (st.release.sys by src, ld.acquire.sys here)
//   flags[FLAG_DONE  + src]  : "src has finished reading my buffer for epoch e"
// Flags only ever grow, so no reset race exists; the host passes the expected epoch.
#include "common.cuh"
#include "symm_common.cuh"

namespace {

template <typename T>
__device__ __forceinline__ void accum8(float (&acc)[16 / sizeof(T)], uint4 raw)
{
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int k = 0; k < (int)(16 / sizeof(T)); k++) acc[k] += dsb::to_f<T>(e[k]);
}


// write VEC results (optionally accumulating into the destination) with 16-byte accesses
template <typename OutT, int VEC>
__device__ __forceinline__ float store_out(OutT* __restrict__ dst, const float (&v)[VEC], int accumulate, float scale)
{
    constexpr int PER = 16 / sizeof(OutT);
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < VEC / PER; c++) {
        uint4 raw;
        OutT* e = reinterpret_cast<OutT*>(&raw);
        if (accumulate) raw = *reinterpret_cast<const uint4*>(dst + c * PER);
#pragma unroll
        for (int k = 0; k < PER; k++) {
            float x = v[c * PER + k] * scale;
            if (accumulate) x += dsb::to_f<OutT>(e[k]);
            ss = fmaf(x, x, ss);
            e[k] = dsb::from_f<OutT>(x);
        }
        *reinterpret_cast<uint4*>(dst + c * PER) = raw;
    }
    return ss;
}

// ---------------------------------------------------------------------------------------------------------
// reduce-scatter: out_shard (+)= scale * sum_p peer_full[p][rank*shard_numel + i]
// `ticket` is a zero-initialised local uint32 used to detect the last CTA (which publishes DONE to peers).
// ---------------------------------------------------------------------------------------------------------
template <typename T, typename OutT>
__global__ void __launch_bounds__(512) symm_reduce_scatter_kernel(PeerPtrs bufs, PeerPtrs flags, int rank, int world,
                                                                  int64_t shard_numel, OutT* __restrict__ out,
                                                                  int accumulate, float scale,
                                                                  float* __restrict__ sumsq_partials,
                                                                  unsigned int* __restrict__ ticket, int ch_ready,
                                                                  int ch_done, uint32_t epoch)
{
    __shared__ float red[33];
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);

    constexpr int VEC = 16 / sizeof(T);
    constexpr int U = 8;  // independent 16-byte loads in flight per thread and peer (NVLink latency ~2 us)
    const int64_t nvec = shard_numel / VEC;
    const int64_t base = (int64_t)rank * shard_numel;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float ss = 0.f;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
        float acc[U][VEC];
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int k = 0; k < VEC; k++) acc[u][k] = 0.f;
        // walk the peers in a rank-rotated order to spread NVLink traffic; 2 peers x U loads issued per round
#pragma unroll 2
        for (int s = 0; s < world; s++) {
            const int p = (rank + s) % world;
            const T* src = reinterpret_cast<const T*>(bufs.p[p]) + base;
            uint4 raw[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                raw[u] = (i < nvec) ? ld_peer_v4(src + i * VEC) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < U; u++) accum8<T>(acc[u], raw[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t i = i0 + u * stride;
            if (i < nvec) ss += store_out<OutT, VEC>(out + i * VEC, acc[u], accumulate, scale);
        }
    }
    if (sumsq_partials != nullptr) {
        ss = dsb::block_sum(ss, red);
        if (threadIdx.x == 0) sumsq_partials[blockIdx.x] = ss;
    }
    // last CTA tells the peers that their buffers are no longer being read by me
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_done, epoch);
}

// multimem variant: the NVSwitch performs the reduction (one load instruction per 16 bytes, fp32 accumulation)
template <typename OutT>
__global__ void __launch_bounds__(512) symm_reduce_scatter_mc_bf16_kernel(const void* mc_base, PeerPtrs flags, int rank,
                                                                          int world, int64_t shard_numel,
                                                                          OutT* __restrict__ out, int accumulate,
                                                                          float scale, float* __restrict__ sumsq_partials,
                                                                          unsigned int* __restrict__ ticket,
                                                                          int ch_ready, int ch_done, uint32_t epoch)
{
    __shared__ float red[33];
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);
    const int64_t nvec = shard_numel / 8;
    constexpr int U = 8;
    const __nv_bfloat16* src0 = reinterpret_cast<const __nv_bfloat16*>(mc_base) + (int64_t)rank * shard_numel;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    float ss = 0.f;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
        uint4 r[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t i = i0 + u * stride;
            r[u] = make_uint4(0, 0, 0, 0);
            if (i < nvec)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                             : "=r"(r[u].x), "=r"(r[u].y), "=r"(r[u].z), "=r"(r[u].w)
                             : "l"(src0 + i * 8)
                             : "memory");
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t i = i0 + u * stride;
            if (i >= nvec) continue;
            const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&r[u]);
            float acc[8];
#pragma unroll
            for (int k = 0; k < 8; k++) acc[k] = __bfloat162float(e[k]);
            ss += store_out<OutT, 8>(out + i * 8, acc, accumulate, scale);
        }
    }
    if (sumsq_partials != nullptr) {
        ss = dsb::block_sum(ss, red);
        if (threadIdx.x == 0) sumsq_partials[blockIdx.x] = ss;
    }
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_done, epoch);
}

// ---------------------------------------------------------------------------------------------------------
// all-gather (pull): full[p*shard + i] = peer_shard[p][i] for every p. 16-byte granularity, any dtype.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) symm_all_gather_kernel(PeerPtrs shards, PeerPtrs flags, int rank, int world,
                                                              int64_t shard_bytes, uint8_t* __restrict__ full,
                                                              unsigned int* __restrict__ ticket, int ch_ready,
                                                              int ch_done, uint32_t epoch)
{
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);
    const int64_t nvec = shard_bytes / 16;
    constexpr int U = 16;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int s = 0; s < world; s++) {
        const int p = (rank + s) % world;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(shards.p[p]);
        uint8_t* dst = full + (int64_t)p * shard_bytes;
        for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < nvec) v[u] = ld_peer_v4(src + i * 16);
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < nvec) *reinterpret_cast<uint4*>(dst + i * 16) = v[u];
            }
        }
    }
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_done, epoch);
}

// all-gather (multicast push): every rank stores its shard once to the multicast address of the symmetric full
// buffer; the switch replicates it into all peers' buffers. Completion: READY flags after a system fence.
__global__ void __launch_bounds__(512) symm_all_gather_mc_kernel(const uint8_t* __restrict__ my_shard, void* mc_full,
                                                                 PeerPtrs flags, int rank, int world,
                                                                 int64_t shard_bytes, unsigned int* __restrict__ ticket,
                                                                 int ch_ready, int ch_done, uint32_t epoch,
                                                                 uint32_t done_epoch)
{
    // the destination buffers of ALL peers must have been released by their consumers (DONE >= done_epoch)
    if (done_epoch != 0) wait_all(flags, rank, world, ch_done, done_epoch);
    const int64_t nvec = shard_bytes / 16;
    uint8_t* dst0 = reinterpret_cast<uint8_t*>(mc_full) + (int64_t)rank * shard_bytes;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        const uint4 v = *reinterpret_cast<const uint4*>(my_shard + i * 16);
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst0 + i * 16), "r"(v.x),
                     "r"(v.y), "r"(v.z), "r"(v.w)
                     : "memory");
    }
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1);
    __syncthreads();
    if (last) {
        __threadfence_system();
        signal_all(flags, rank, world, ch_ready, epoch);
        wait_all(flags, rank, world, ch_ready, epoch);  // everyone's shard has landed in my buffer
    }
}

// ---------------------------------------------------------------------------------------------------------
// all-to-all (pull): out[p-th chunk] = peer_in[p][rank-th chunk]; optional 2-level layout so the Ulysses
// [P, s/P, h/P, d] <-> head/sequence re-shard needs no separate permute pass:
//   peer_in[p] is viewed as [world][rows][row_bytes]; out as [rows][world][row_bytes] when `interleave` else
//   [world][rows][row_bytes].
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) symm_all_to_all_kernel(PeerPtrs ins, PeerPtrs flags, int rank, int world,
                                                              int64_t rows, int64_t row_bytes, int interleave,
                                                              uint8_t* __restrict__ out,
                                                              unsigned int* __restrict__ ticket, int ch_ready,
                                                              int ch_done, uint32_t epoch)
{
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);
    const int64_t vec_per_row = row_bytes / 16;
    const int64_t per_peer = rows * vec_per_row;
    constexpr int U = 16;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int s = 0; s < world; s++) {
        const int p = (rank + s) % world;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(ins.p[p]) + (int64_t)rank * rows * row_bytes;
        for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < per_peer; i0 += stride * U) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < per_peer) v[u] = ld_peer_v4(src + i * 16);
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < per_peer) {
                    const int64_t r = i / vec_per_row, c = i - r * vec_per_row;
                    uint8_t* dst = interleave ? out + (r * world + p) * row_bytes + c * 16
                                              : out + ((int64_t)p * rows + r) * row_bytes + c * 16;
                    *reinterpret_cast<uint4*>(dst) = v[u];
                }
            }
        }
    }
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_done, epoch);
}

// Spin (one CTA) until all peers reached `epoch` on `channel`: inserted on the producer stream before a
// symmetric buffer is overwritten.
__global__ void symm_wait_kernel(PeerPtrs flags, int rank, int world, int channel, uint32_t epoch)
{
    wait_all(flags, rank, world, channel, epoch);
}

__global__ void symm_signal_kernel(PeerPtrs flags, int rank, int world, int channel, uint32_t epoch)
{
    __threadfence_system();
    signal_all(flags, rank, world, channel, epoch);
}

PeerPtrs make_ptrs(const uint64_t* host_ptrs, int world)
{
    PeerPtrs pp;
    for (int i = 0; i < MAX_WORLD; i++) pp.p[i] = (i < world) ? reinterpret_cast<void*>(host_ptrs[i]) : nullptr;
    return pp;
}

int comm_ctas(int64_t bytes, int requested)
{
    if (requested > 0) return requested;
    int64_t c = bytes / (512 * 16 * 8);
    if (c < 1) c = 1;
    if (c > 64) c = 64;  // leave the remaining SMs to the overlapped GEMMs
    return (int)c;
}

}  // namespace

extern "C" {

// in_dt: element type of the peer buffers (1 fp16, 2 bf16, 0 fp32); out_dt: type of the local shard accumulator.
// mc_ptr != 0 selects the multimem (NVSwitch in-switch reduction) path for bf16 inputs.
int dsb200_symm_reduce_scatter(const uint64_t* peer_bufs, const uint64_t* peer_flags, uint64_t mc_ptr, int rank,
                               int world, int64_t shard_numel, int in_dt, void* out, int out_dt, int accumulate,
                               float scale, float* sumsq_partials, unsigned int* ticket, int ch_ready, int ch_done,
                               unsigned int epoch, int ctas, cudaStream_t stream)
{
    if (world > MAX_WORLD) {
        dsb200_set_error("symm: world size > 16");
        return -1;
    }
    const int esz = in_dt == DSB_DT_F32 ? 4 : 2;
    if ((shard_numel * esz) % 16 != 0) {
        dsb200_set_error("symm_reduce_scatter: shard must be a multiple of 16 bytes");
        return -1;
    }
    PeerPtrs bufs = make_ptrs(peer_bufs, world), flags = make_ptrs(peer_flags, world);
    const int grid = comm_ctas(shard_numel * esz * world, ctas);
#define RS_LAUNCH(T, O)                                                                                           \
    symm_reduce_scatter_kernel<T, O><<<grid, 512, 0, stream>>>(bufs, flags, rank, world, shard_numel, (O*)out,    \
                                                               accumulate, scale, sumsq_partials, ticket, ch_ready, \
                                                               ch_done, epoch)
    if (mc_ptr != 0 && in_dt == DSB_DT_BF16) {
        if (out_dt == DSB_DT_F32)
            symm_reduce_scatter_mc_bf16_kernel<float><<<grid, 512, 0, stream>>>(
                (const void*)mc_ptr, flags, rank, world, shard_numel, (float*)out, accumulate, scale, sumsq_partials,
                ticket, ch_ready, ch_done, epoch);
        else
            symm_reduce_scatter_mc_bf16_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(
                (const void*)mc_ptr, flags, rank, world, shard_numel, (__nv_bfloat16*)out, accumulate, scale,
                sumsq_partials, ticket, ch_ready, ch_done, epoch);
    } else if (in_dt == DSB_DT_BF16 && out_dt == DSB_DT_BF16) {
        RS_LAUNCH(__nv_bfloat16, __nv_bfloat16);
    } else if (in_dt == DSB_DT_BF16 && out_dt == DSB_DT_F32) {
        RS_LAUNCH(__nv_bfloat16, float);
    } else if (in_dt == DSB_DT_F16 && out_dt == DSB_DT_F16) {
        RS_LAUNCH(__half, __half);
    } else if (in_dt == DSB_DT_F16 && out_dt == DSB_DT_F32) {
        RS_LAUNCH(__half, float);
    } else if (in_dt == DSB_DT_F32 && out_dt == DSB_DT_F32) {
        RS_LAUNCH(float, float);
    } else {
        dsb200_set_error("symm_reduce_scatter: unsupported dtype combination");
        return -1;
    }
#undef RS_LAUNCH
    DSB_CHECK_LAUNCH("symm_reduce_scatter");
    return grid;
}

int dsb200_symm_all_gather(const uint64_t* peer_shards, const uint64_t* peer_flags, uint64_t mc_full, int rank,
                           int world, int64_t shard_bytes, const void* my_shard, void* full, unsigned int* ticket,
                           int ch_ready, int ch_done, unsigned int epoch, unsigned int done_epoch, int ctas,
                           cudaStream_t stream)
{
    if (world > MAX_WORLD || shard_bytes % 16 != 0) {
        dsb200_set_error("symm_all_gather: world > 16 or shard not a multiple of 16 bytes");
        return -1;
    }
    PeerPtrs flags = make_ptrs(peer_flags, world);
    const int grid = comm_ctas(shard_bytes * world, ctas);
    if (mc_full != 0) {
        symm_all_gather_mc_kernel<<<grid, 512, 0, stream>>>((const uint8_t*)my_shard, (void*)mc_full, flags, rank, world,
                                                            shard_bytes, ticket, ch_ready, ch_done, epoch, done_epoch);
    } else {
        PeerPtrs shards = make_ptrs(peer_shards, world);
        symm_all_gather_kernel<<<grid, 512, 0, stream>>>(shards, flags, rank, world, shard_bytes, (uint8_t*)full, ticket,
                                                         ch_ready, ch_done, epoch);
    }
    DSB_CHECK_LAUNCH("symm_all_gather");
    return 0;
}

int dsb200_symm_all_to_all(const uint64_t* peer_ins, const uint64_t* peer_flags, int rank, int world, int64_t rows,
                           int64_t row_bytes, int interleave, void* out, unsigned int* ticket, int ch_ready,
                           int ch_done, unsigned int epoch, int ctas, cudaStream_t stream)
{
    if (world > MAX_WORLD || row_bytes % 16 != 0) {
        dsb200_set_error("symm_all_to_all: world > 16 or row not a multiple of 16 bytes");
        return -1;
    }
    PeerPtrs ins = make_ptrs(peer_ins, world), flags = make_ptrs(peer_flags, world);
    const int grid = comm_ctas(rows * row_bytes * world, ctas);
    symm_all_to_all_kernel<<<grid, 512, 0, stream>>>(ins, flags, rank, world, rows, row_bytes, interleave, (uint8_t*)out,
                                                     ticket, ch_ready, ch_done, epoch);
    DSB_CHECK_LAUNCH("symm_all_to_all");
    return 0;
}

// All-gather on the copy engines: the peers' shards are pulled with cudaMemcpyAsync over NVLink (DMA, no SM is
// occupied, so the overlapped GEMMs keep all 148 SMs), spread over `n_helpers` helper streams so that several copy
// engines run at once; `stream` forks before and joins after. No flag handshake here: the caller orders the gather
// against the writers of the shards (symm_signal after the optimizer step / symm_wait before the first gather) and
// publishes completion with symm_signal when the shards are about to be rewritten.
// Reference counterpart: the DeepCompile executor's symmetric-memory all-gather (csrc/compile/z3.cpp:100-112), which
// brackets per-peer copy_() calls with two barriers for every parameter.
int dsb200_symm_all_gather_ce(const uint64_t* peer_shards, int rank, int world, int64_t shard_bytes, void* full,
                              const cudaStream_t* helpers, int n_helpers, cudaStream_t stream)
{
    constexpr int MAX_HELPERS = 8;
    if (world > MAX_WORLD || world < 1 || n_helpers > MAX_HELPERS || n_helpers < 0) {
        dsb200_set_error("symm_all_gather_ce: world > 16 or more than 8 helper streams");
        return -1;
    }
    struct Events {
        cudaEvent_t fork = nullptr, join[MAX_HELPERS] = {};
    };
    static Events per_dev[64];
    int dev = 0;
    DSB_CHECK(cudaGetDevice(&dev));
    Events& ev = per_dev[dev & 63];
    if (ev.fork == nullptr) {
        DSB_CHECK(cudaEventCreateWithFlags(&ev.fork, cudaEventDisableTiming));
        for (int i = 0; i < MAX_HELPERS; i++) DSB_CHECK(cudaEventCreateWithFlags(&ev.join[i], cudaEventDisableTiming));
    }
    uint8_t* dst = reinterpret_cast<uint8_t*>(full);
    const int used = n_helpers < world - 1 ? n_helpers : world - 1;
    if (used > 0) {
        DSB_CHECK(cudaEventRecord(ev.fork, stream));
        for (int h = 0; h < used; h++) DSB_CHECK(cudaStreamWaitEvent(helpers[h], ev.fork, 0));
    }
    for (int i = 1; i < world; i++) {
        const int p = (rank + i) % world;  // every rank starts with a different peer: no hot spot on one GPU
        cudaStream_t s = used > 0 ? helpers[(i - 1) % used] : stream;
        DSB_CHECK(cudaMemcpyAsync(dst + (int64_t)p * shard_bytes, reinterpret_cast<const void*>(peer_shards[p]),
                                  (size_t)shard_bytes, cudaMemcpyDeviceToDevice, s));
    }
    DSB_CHECK(cudaMemcpyAsync(dst + (int64_t)rank * shard_bytes, reinterpret_cast<const void*>(peer_shards[rank]),
                              (size_t)shard_bytes, cudaMemcpyDeviceToDevice, stream));
    for (int h = 0; h < used; h++) {
        DSB_CHECK(cudaEventRecord(ev.join[h], helpers[h]));
        DSB_CHECK(cudaStreamWaitEvent(stream, ev.join[h], 0));
    }
    return 0;
}

int dsb200_symm_wait(const uint64_t* peer_flags, int rank, int world, int channel, unsigned int epoch,
                     cudaStream_t stream)
{
    PeerPtrs flags = make_ptrs(peer_flags, world);
    symm_wait_kernel<<<1, 32, 0, stream>>>(flags, rank, world, channel, epoch);
    DSB_CHECK_LAUNCH("symm_wait");
    return 0;
}

int dsb200_symm_signal(const uint64_t* peer_flags, int rank, int world, int channel, unsigned int epoch,
                       cudaStream_t stream)
{
    PeerPtrs flags = make_ptrs(peer_flags, world);
    symm_signal_kernel<<<1, 32, 0, stream>>>(flags, rank, world, channel, epoch);
    DSB_CHECK_LAUNCH("symm_signal");
    return 0;
}

}  // extern "C"
