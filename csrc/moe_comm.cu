// Expert-parallel MoE dispatch / combine as single kernels over NVLink peer memory (symmetric buffers).
//
//   dispatch   the expert rank PULLS the token rows routed to its experts straight out of every peer's activation
//              buffer (indexed by that peer's expert-sorted token list) and lands them expert-contiguous: the
//              sender-side gather, the all-to-all-v and the receiver-side permute of the reference are one pass, and no
//              row is written twice. An optional per-slot scale (router weight) serves the backward of combine.
//   combine    the token owner PULLS the k expert outputs of each of its tokens from the expert ranks' output buffers,
//              applies the router weights, sums in fp32 and writes the token row once: the un-permute, the return
//              all-to-all-v, the scale and the index_add of the reference are one pass (deterministic, no atomics).
//              With unit weights it is the backward of dispatch. The raw pulled rows can be kept for the backward.
//
// Reference: module_inject/auto_ep_layer.py:184-350 (_AllToAllV dispatch / combine autograd functions), :575-730
// (router -> sort -> all-to-all-v -> permute -> experts -> unpermute -> all-to-all-v -> weighted combine),
// moe/ep_kernels.py:72-310 (Triton permutation index kernel); moe/sharded_moe.py:104 (_AllToAll).
// Ordering: epoch flags of symm_common.cuh (READY: "my staging buffers are final", DONE: "I finished reading yours").
#include "symm_common.cuh"

namespace {

struct MoeSeg {
    int src_rank;   // peer the rows come from
    int src_start;  // first slot in that peer's expert-sorted list
    int dst_start;  // first row of the local (expert-contiguous) output
    int n;          // rows
};

__device__ __forceinline__ int ld_peer_i32(const void* p)
{
    int v;
    asm volatile("ld.global.relaxed.sys.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ld_peer_f32(const void* p)
{
    float v;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ uint4 scale_bf16x8(uint4 v, float s)
{
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float2 f = __bfloat1622float2(h[i]);
        h[i] = __floats2bfloat162_rn(f.x * s, f.y * s);
    }
    return v;
}

// One warp per destination row. segs are sorted by dst_start and cover [0, n_rows) without gaps.
__global__ void __launch_bounds__(512) moe_dispatch_kernel(PeerPtrs xs, PeerPtrs idxs, PeerPtrs scales, PeerPtrs flags,
                                                           int rank, int world, const MoeSeg* __restrict__ segs, int nseg,
                                                           int n_rows, int row_bytes, uint8_t* __restrict__ out,
                                                           unsigned int* __restrict__ ticket, int ch_ready, int ch_done,
                                                           uint32_t epoch)
{
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);
    const int warps_per_cta = blockDim.x >> 5, lane = threadIdx.x & 31;
    const int vec_per_row = row_bytes >> 4;
    const bool has_scale = scales.p[0] != nullptr;
    for (int r = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); r < n_rows; r += gridDim.x * warps_per_cta) {
        // segment of this row (binary search on dst_start)
        int lo = 0, hi = nseg - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (segs[mid].dst_start <= r) lo = mid;
            else hi = mid - 1;
        }
        const MoeSeg sg = segs[lo];
        const int slot = sg.src_start + (r - sg.dst_start);
        const int tok = ld_peer_i32(reinterpret_cast<const int*>(idxs.p[sg.src_rank]) + slot);
        const float sc = has_scale ? ld_peer_f32(reinterpret_cast<const float*>(scales.p[sg.src_rank]) + slot) : 1.f;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(xs.p[sg.src_rank]) + (int64_t)tok * row_bytes;
        uint8_t* dst = out + (int64_t)r * row_bytes;
        for (int c0 = lane; c0 < vec_per_row; c0 += 32 * 4) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (c0 + u * 32 < vec_per_row) v[u] = ld_peer_v4(src + (int64_t)(c0 + u * 32) * 16);
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (c0 + u * 32 < vec_per_row)
                    *reinterpret_cast<uint4*>(dst + (int64_t)(c0 + u * 32) * 16) = has_scale ? scale_bf16x8(v[u], sc) : v[u];
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

// One warp per token: out[t] = sum_j w[t, j] * peer_y[slot_rank[t, j]][slot_row[t, j]].
template <int K>
__global__ void __launch_bounds__(512) moe_combine_kernel(PeerPtrs ys, PeerPtrs flags, int rank, int world,
                                                          const int* __restrict__ slot_rank,
                                                          const int* __restrict__ slot_row,
                                                          const float* __restrict__ weights, int T, int row_bytes,
                                                          uint8_t* __restrict__ out, uint8_t* __restrict__ rows_out,
                                                          unsigned int* __restrict__ ticket, int ch_ready, int ch_done,
                                                          uint32_t epoch)
{
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);
    const int warps_per_cta = blockDim.x >> 5, lane = threadIdx.x & 31;
    const int vec_per_row = row_bytes >> 4;
    for (int t = blockIdx.x * warps_per_cta + (threadIdx.x >> 5); t < T; t += gridDim.x * warps_per_cta) {
        const uint8_t* src[K];
        float w[K];
#pragma unroll
        for (int j = 0; j < K; j++) {
            const int sr = slot_rank[t * K + j], row = slot_row[t * K + j];
            src[j] = (sr >= 0) ? reinterpret_cast<const uint8_t*>(ys.p[sr]) + (int64_t)row * row_bytes : nullptr;
            w[j] = weights != nullptr ? weights[t * K + j] : 1.f;
        }
        for (int c = lane; c < vec_per_row; c += 32) {
            uint4 v[K];
#pragma unroll
            for (int j = 0; j < K; j++) v[j] = src[j] != nullptr ? ld_peer_v4(src[j] + (int64_t)c * 16) : make_uint4(0, 0, 0, 0);
            float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int j = 0; j < K; j++) {
                const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v[j]);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const float2 f = __bfloat1622float2(h[i]);
                    acc[2 * i] = fmaf(w[j], f.x, acc[2 * i]);
                    acc[2 * i + 1] = fmaf(w[j], f.y, acc[2 * i + 1]);
                }
                if (rows_out != nullptr)
                    *reinterpret_cast<uint4*>(rows_out + ((int64_t)t * K + j) * row_bytes + (int64_t)c * 16) = v[j];
            }
            uint4 o;
            __nv_bfloat162* oh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
            for (int i = 0; i < 4; i++) oh[i] = __floats2bfloat162_rn(acc[2 * i], acc[2 * i + 1]);
            *reinterpret_cast<uint4*>(out + (int64_t)t * row_bytes + (int64_t)c * 16) = o;
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

int moe_ctas(int64_t rows, int requested)
{
    if (requested > 0) return requested;
    int64_t c = (rows + 15) / 16;  // 16 warps per CTA
    if (c < 1) c = 1;
    if (c > 64) c = 64;
    return (int)c;
}

}  // namespace

extern "C" {

// peer_x[p]: base of rank p's activation staging buffer [tokens, H] bf16; peer_idx[p]: its expert-sorted token list
// (int32); peer_scale[p]: per-slot fp32 scale or all zero pointers; segs: device array of nseg MoeSeg (16 bytes each)
// sorted by dst_start, covering [0, n_rows). out [n_rows, H] bf16. Returns the grid size or < 0.
int dsb200_moe_dispatch(const uint64_t* peer_x, const uint64_t* peer_idx, const uint64_t* peer_scale,
                        const uint64_t* peer_flags, int rank, int world, const void* segs, int nseg, int n_rows,
                        int hidden, void* out, unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch,
                        int ctas, cudaStream_t stream)
{
    if (world > MAX_WORLD || hidden % 8 != 0) {
        dsb200_set_error("moe_dispatch: world > 16 or hidden size not a multiple of 8");
        return -1;
    }
    PeerPtrs xs = make_peer_ptrs(peer_x, world), idxs = make_peer_ptrs(peer_idx, world);
    PeerPtrs flags = make_peer_ptrs(peer_flags, world);
    PeerPtrs sc = make_peer_ptrs(peer_scale, peer_scale != nullptr && peer_scale[0] != 0 ? world : 0);
    const int grid = moe_ctas(n_rows, ctas);
    moe_dispatch_kernel<<<grid, 512, 0, stream>>>(xs, idxs, sc, flags, rank, world, (const MoeSeg*)segs, nseg, n_rows,
                                                  hidden * 2, (uint8_t*)out, ticket, ch_ready, ch_done, epoch);
    DSB_CHECK_LAUNCH("moe_dispatch");
    return grid;
}

// peer_y[p]: base of rank p's expert-output buffer [rows, H] bf16. slot_rank / slot_row: int32 [T * k] (rank < 0:
// dropped slot, contributes zero); weights fp32 [T * k] or null (all ones); out [T, H] bf16; rows_out [T * k, H] bf16
// or null. k in {1, 2, 4, 6, 8}.
int dsb200_moe_combine(const uint64_t* peer_y, const uint64_t* peer_flags, int rank, int world, const int* slot_rank,
                       const int* slot_row, const float* weights, int T, int k, int hidden, void* out, void* rows_out,
                       unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch, int ctas,
                       cudaStream_t stream)
{
    if (world > MAX_WORLD || hidden % 8 != 0) {
        dsb200_set_error("moe_combine: world > 16 or hidden size not a multiple of 8");
        return -1;
    }
    PeerPtrs ys = make_peer_ptrs(peer_y, world), flags = make_peer_ptrs(peer_flags, world);
    const int grid = moe_ctas(T, ctas);
#define MC_LAUNCH(KK)                                                                                                 \
    moe_combine_kernel<KK><<<grid, 512, 0, stream>>>(ys, flags, rank, world, slot_rank, slot_row, weights, T, hidden * 2, \
                                                     (uint8_t*)out, (uint8_t*)rows_out, ticket, ch_ready, ch_done, epoch)
    switch (k) {
        case 1: MC_LAUNCH(1); break;
        case 2: MC_LAUNCH(2); break;
        case 4: MC_LAUNCH(4); break;
        case 6: MC_LAUNCH(6); break;
        case 8: MC_LAUNCH(8); break;
        default: dsb200_set_error("moe_combine: top-k must be 1, 2, 4, 6 or 8"); return -1;
    }
#undef MC_LAUNCH
    DSB_CHECK_LAUNCH("moe_combine");
    return grid;
}

}  // extern "C"
