// Fused collective + compute kernels over symmetric memory (second generation; the plain collectives are in symm.cu).
//
//  (A) reduce-scatter + unscale/average + partitioned AdamW update + bf16 weight write   -- north-star path (1)
//      Reference: reduce_scatter_coalesced (runtime/comm/coalesced_collectives.py:158) then, at step time, a separate
//      FusedAdam launch per sub-group (ops/adam/fused_adam.py:195, stage3.py:1072 step). Here the rank pulls its
//      gradient shard from every peer over NVLink, sums in fp32 registers and applies the optimizer to the fp32
//      master/moments in the same pass: the reduced gradient never touches HBM. Valid when the step does not need
//      a global gradient norm first (no clipping, bf16 = no loss scale) and gradient accumulation is 1.
//  (B) qwZ all-gather: block-scaled FP8 quantize -> pull over NVLink -> dequantize      -- north-star path (3)
//      Reference: zero_quantized_weights (partition_parameters.py:1248 all_gather_coalesced + CUDAQuantizer): three
//      kernels and two NCCL calls (payload + scales). Here ONE kernel quantizes the local bf16 shard into a
//      symmetric staging buffer, publishes it with an epoch flag, then pulls the peers' FP8 payload (half the
//      NVLink bytes) and scales and writes dequantized bf16 straight into the gathered parameter buffer.
#include "symm_common.cuh"
#include <cuda_fp8.h>

namespace {

constexpr int MAX_SEG = 8;

struct FusedAdamArgs {
    int nseg;
    int64_t bound[MAX_SEG + 1];  // element offsets inside the shard; segment s = [bound[s], bound[s+1])
    float h[MAX_SEG][8];         // lr, beta1, beta2, eps, wd, bias_correction1, bias_correction2, mode(1 = AdamW)
};

__device__ __forceinline__ void adam1(float& p, float g, float& m, float& v, const float* h)
{
    const float lr = h[0], b1 = h[1], b2 = h[2], eps = h[3], wd = h[4], bc1 = h[5], bc2 = h[6];
    const int mode = (int)h[7];
    if (mode == 0) g = fmaf(wd, p, g);
    m = fmaf(b1, m, (1.f - b1) * g);
    v = fmaf(b2, v, (1.f - b2) * g * g);
    float upd = (m / bc1) / (sqrtf(v / bc2) + eps);
    if (mode == 1) upd = fmaf(wd, p, upd);
    p = fmaf(-lr, upd, p);
}

template <typename T>
__device__ __forceinline__ void accum_vec(float (&acc)[16 / sizeof(T)], uint4 raw)
{
    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int k = 0; k < (int)(16 / sizeof(T)); k++) acc[k] += dsb::to_f<T>(e[k]);
}

// ---------------------------------------------------------------------------------------------------------
// (A) reduce-scatter + AdamW. bufs[p] = peer p's full gradient bucket (world * shard_numel elements of T).
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(512) symm_rs_adam_kernel(PeerPtrs bufs, PeerPtrs flags, int rank, int world,
                                                           int64_t shard_numel, float gscale, float* __restrict__ P,
                                                           float* __restrict__ M, float* __restrict__ V,
                                                           __nv_bfloat16* __restrict__ LP, FusedAdamArgs a,
                                                           float* __restrict__ sumsq_partials,
                                                           unsigned int* __restrict__ ticket, int ch_ready, int ch_done,
                                                           uint32_t epoch)
{
    __shared__ float red[33];
    if (blockIdx.x == 0) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);

    constexpr int VEC = 16 / sizeof(T);
    constexpr int U = 4;
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
            for (int u = 0; u < U; u++) accum_vec<T>(acc[u], raw[u]);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int64_t i = i0 + u * stride;
            if (i >= nvec) continue;
            const int64_t e0 = i * VEC;
            int seg = 0;
            while (seg + 1 < a.nseg && e0 >= a.bound[seg + 1]) seg++;
            if (e0 >= a.bound[a.nseg]) continue;  // padding behind the last parameter: nothing to update
            const float* h = a.h[seg];
#pragma unroll
            for (int c = 0; c < VEC / 4; c++) {
                float4 p = *reinterpret_cast<const float4*>(P + e0 + 4 * c);
                float4 m = *reinterpret_cast<const float4*>(M + e0 + 4 * c);
                float4 v = *reinterpret_cast<const float4*>(V + e0 + 4 * c);
                float g0 = acc[u][4 * c] * gscale, g1 = acc[u][4 * c + 1] * gscale;
                float g2 = acc[u][4 * c + 2] * gscale, g3 = acc[u][4 * c + 3] * gscale;
                ss = fmaf(g0, g0, fmaf(g1, g1, fmaf(g2, g2, fmaf(g3, g3, ss))));
                adam1(p.x, g0, m.x, v.x, h);
                adam1(p.y, g1, m.y, v.y, h);
                adam1(p.z, g2, m.z, v.z, h);
                adam1(p.w, g3, m.w, v.w, h);
                *reinterpret_cast<float4*>(P + e0 + 4 * c) = p;
                *reinterpret_cast<float4*>(M + e0 + 4 * c) = m;
                *reinterpret_cast<float4*>(V + e0 + 4 * c) = v;
                if (LP != nullptr) {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(p.x, p.y), hi = __floats2bfloat162_rn(p.z, p.w);
                    uint2 o;
                    o.x = *reinterpret_cast<uint32_t*>(&lo);
                    o.y = *reinterpret_cast<uint32_t*>(&hi);
                    *reinterpret_cast<uint2*>(LP + e0 + 4 * c) = o;
                }
            }
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
// (B) FP8 (e4m3, one fp32 scale per `block` elements) all-gather. stage[p] = peer p's symmetric staging buffer:
// [ shard_elems bytes of payload | pad to 256 | shard_elems / block fp32 scales ].
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint16_t to_e4m3x2(float lo, float hi)
{
    uint16_t r;
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float2 from_e4m3x2(uint16_t v)
{
    uint32_t h2;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(v));
    return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
__device__ __forceinline__ float ld_peer_f32(const float* p)
{
    float r;
    asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
    return r;
}

template <typename T>
__device__ __forceinline__ void dequant16_store(T* dst, uint4 raw, float s)
{
    const uint16_t* q = reinterpret_cast<const uint16_t*>(&raw);
    dsb::f8 lo, hi;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const float2 a = from_e4m3x2(q[k]), b = from_e4m3x2(q[4 + k]);
        lo.v[2 * k] = a.x * s, lo.v[2 * k + 1] = a.y * s;
        hi.v[2 * k] = b.x * s, hi.v[2 * k + 1] = b.y * s;
    }
    dsb::store8(dst, lo);
    dsb::store8(dst + 8, hi);
}

template <typename T>
__global__ void __launch_bounds__(512) symm_all_gather_fp8_kernel(const T* __restrict__ my_shard, PeerPtrs stage,
                                                                  PeerPtrs flags, int rank, int world,
                                                                  int64_t shard_elems, int block, int64_t scale_off,
                                                                  T* __restrict__ full,
                                                                  unsigned int* __restrict__ tickets, int ch_ready,
                                                                  int ch_done, uint32_t epoch, uint32_t reuse_epoch)
{
    // my staging buffer may still be read by peers working on the collective that used it last
    if (reuse_epoch != 0) wait_all(flags, rank, world, ch_done, reuse_epoch);

    // ---- phase 1: quantize my shard (one warp per block), keep the dequantized values as my slice of `full` so
    //      that every rank computes with bit-identical weights
    uint8_t* myq = reinterpret_cast<uint8_t*>(stage.p[rank]);
    float* myscale = reinterpret_cast<float*>(myq + scale_off);
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nblocks = shard_elems / block;
    const int iters = block / 256;
    T* mine = full + (int64_t)rank * shard_elems;
    for (int64_t b = warp; b < nblocks; b += nwarps) {
        const T* xb = my_shard + b * block;
        float amax = 0.f;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xb + (it * 32 + lane) * 8);
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v.v[i]));
        }
        amax = dsb::warp_max(amax);
        const float scale = amax > 0.f ? amax / 448.f : 1.f;
        const float inv = 1.f / scale;
        if (lane == 0) myscale[b] = scale;
        for (int it = 0; it < iters; it++) {
            const int64_t off = b * block + (it * 32 + lane) * 8;
            dsb::f8 v = dsb::load8(my_shard + off);
            uint16_t p[4];
            dsb::f8 dq;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                p[i] = to_e4m3x2(v.v[2 * i] * inv, v.v[2 * i + 1] * inv);
                const float2 f = from_e4m3x2(p[i]);
                dq.v[2 * i] = f.x * scale, dq.v[2 * i + 1] = f.y * scale;
            }
            uint2 out;
            out.x = (uint32_t)p[0] | ((uint32_t)p[1] << 16);
            out.y = (uint32_t)p[2] | ((uint32_t)p[3] << 16);
            *reinterpret_cast<uint2*>(myq + off) = out;
            dsb::store8(mine + off, dq);
        }
    }
    // ---- publish: the last CTA to finish phase 1 raises READY (which doubles as the grid barrier, since every
    //      CTA waits for all ranks' READY including its own)
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        last = (atomicInc(tickets, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_ready, epoch);
    wait_all(flags, rank, world, ch_ready, epoch);

    // ---- phase 2: pull FP8 payload + scales from the peers, dequantize into `full`
    constexpr int U = 8;
    const int64_t nvec = shard_elems / 16;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int s = 1; s < world; s++) {
        const int p = (rank + s) % world;
        const uint8_t* q = reinterpret_cast<const uint8_t*>(stage.p[p]);
        const float* sc = reinterpret_cast<const float*>(q + scale_off);
        T* dst = full + (int64_t)p * shard_elems;
        for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
            uint4 raw[U];
            float scl[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < nvec) {
                    raw[u] = ld_peer_v4(q + i * 16);
                    scl[u] = ld_peer_f32(sc + (i * 16) / block);
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int64_t i = i0 + u * stride;
                if (i < nvec) dequant16_store<T>(dst + i * 16, raw[u], scl[u]);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        last = (atomicInc(tickets + 1, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) signal_all(flags, rank, world, ch_done, epoch);
}

int fused_ctas(int64_t bytes, int requested)
{
    if (requested > 0) return requested > DSB_NUM_SMS ? DSB_NUM_SMS : requested;
    int64_t c = bytes / (512 * 16 * 8);
    if (c < 1) c = 1;
    if (c > 64) c = 64;
    return (int)c;
}

}  // namespace

extern "C" {

// Gradient bucket of `world * shard_numel` elements (in_dt 0 fp32 / 2 bf16) in every peer's symmetric buffer ->
// AdamW on my shard. seg_bounds[nseg+1] are element offsets (multiples of 8) inside the shard, hyper[nseg][8] =
// {lr, beta1, beta2, eps, weight_decay, bias_correction1, bias_correction2, mode}. lp may be NULL.
// sumsq_partials (may be NULL) receives one partial of sum(g^2) per CTA; returns the grid size (> 0) or < 0.
int dsb200_symm_rs_adam(const uint64_t* peer_bufs, const uint64_t* peer_flags, int rank, int world,
                        int64_t shard_numel, int in_dt, float grad_scale, float* p, float* m, float* v, void* lp,
                        int nseg, const int64_t* seg_bounds, const float* hyper, float* sumsq_partials,
                        unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch, int ctas,
                        cudaStream_t stream)
{
    if (world > MAX_WORLD || nseg < 1 || nseg > MAX_SEG) {
        dsb200_set_error("symm_rs_adam: world > 16 or segment count not in [1, 8]");
        return -1;
    }
    const int esz = in_dt == DSB_DT_F32 ? 4 : 2;
    const int vec = 16 / esz;
    if (shard_numel % vec != 0 || (in_dt != DSB_DT_F32 && in_dt != DSB_DT_BF16)) {
        dsb200_set_error("symm_rs_adam: shard must be a multiple of 16 bytes of fp32 or bf16");
        return -1;
    }
    FusedAdamArgs a;
    a.nseg = nseg;
    for (int s = 0; s <= nseg; s++) {
        if (seg_bounds[s] % 8 != 0) {
            dsb200_set_error("symm_rs_adam: segment bounds must be multiples of 8 elements");
            return -1;
        }
        a.bound[s] = seg_bounds[s];
    }
    for (int s = 0; s < nseg; s++)
        for (int k = 0; k < 8; k++) a.h[s][k] = hyper[s * 8 + k];
    PeerPtrs bufs = make_peer_ptrs(peer_bufs, world), flags = make_peer_ptrs(peer_flags, world);
    const int grid = fused_ctas(shard_numel * esz * world, ctas);
    if (in_dt == DSB_DT_BF16)
        symm_rs_adam_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(bufs, flags, rank, world, shard_numel, grad_scale, p,
                                                                    m, v, (__nv_bfloat16*)lp, a, sumsq_partials,
                                                                    ticket, ch_ready, ch_done, epoch);
    else
        symm_rs_adam_kernel<float><<<grid, 512, 0, stream>>>(bufs, flags, rank, world, shard_numel, grad_scale, p, m, v,
                                                            (__nv_bfloat16*)lp, a, sumsq_partials, ticket, ch_ready,
                                                            ch_done, epoch);
    DSB_CHECK_LAUNCH("symm_rs_adam");
    return grid;
}

// qwZ all-gather. my_shard: shard_elems elements (dt 1 fp16 / 2 bf16, multiple of `block`, block multiple of 256);
// peer_stage: symmetric staging buffers of at least dsb200_symm_fp8_stage_bytes(shard_elems, block) bytes;
// full: world * shard_elems elements; tickets: two zero-initialised uint32; reuse_epoch: epoch of the previous
// collective that used this staging buffer (0 if none).
int64_t dsb200_symm_fp8_stage_bytes(int64_t shard_elems, int block)
{
    const int64_t q = (shard_elems + 255) / 256 * 256;
    return q + (shard_elems / block) * 4 + 256;
}

int dsb200_symm_all_gather_fp8(const void* my_shard, int dt, const uint64_t* peer_stage, const uint64_t* peer_flags,
                               int rank, int world, int64_t shard_elems, int block, void* full,
                               unsigned int* tickets, int ch_ready, int ch_done, unsigned int epoch,
                               unsigned int reuse_epoch, int ctas, cudaStream_t stream)
{
    if (world > MAX_WORLD || block % 256 != 0 || shard_elems % block != 0) {
        dsb200_set_error("symm_all_gather_fp8: world > 16, block not a multiple of 256, or shard not a multiple of block");
        return -1;
    }
    if (((uintptr_t)my_shard | (uintptr_t)full) & 15) {
        dsb200_set_error("symm_all_gather_fp8: pointers must be 16-byte aligned");
        return -1;
    }
    PeerPtrs stage = make_peer_ptrs(peer_stage, world), flags = make_peer_ptrs(peer_flags, world);
    const int64_t scale_off = (shard_elems + 255) / 256 * 256;
    const int grid = fused_ctas(shard_elems * world, ctas);
    if (dt == DSB_DT_BF16)
        symm_all_gather_fp8_kernel<__nv_bfloat16><<<grid, 512, 0, stream>>>(
            (const __nv_bfloat16*)my_shard, stage, flags, rank, world, shard_elems, block, scale_off,
            (__nv_bfloat16*)full, tickets, ch_ready, ch_done, epoch, reuse_epoch);
    else if (dt == DSB_DT_F16)
        symm_all_gather_fp8_kernel<__half><<<grid, 512, 0, stream>>>((const __half*)my_shard, stage, flags, rank, world,
                                                                    shard_elems, block, scale_off, (__half*)full,
                                                                    tickets, ch_ready, ch_done, epoch, reuse_epoch);
    else {
        dsb200_set_error("symm_all_gather_fp8: dtype must be fp16 or bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("symm_all_gather_fp8");
    return 0;
}

}  // extern "C"
