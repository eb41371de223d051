// Weight-only quantized skinny GEMM for decode: y[M, N] = x[M, K] * dequant(Wq[N, K])^T, M <= 8 tokens, bf16
// activations, weights stored as block-scaled FP8 (e4m3 bytes) or FP4 (two e2m1 codes per byte) with one fp32 scale per
// `group` consecutive elements along K -- the storage of deepspeed.linear.QuantizedParameter / FP_Quantize.
//
// Decode GEMMs are bound by the weight stream, so the weights cross HBM once in their quantized form (1 or 0.5 B per
// element instead of 2) and are expanded in registers with the hardware cvt (e4m3x2 / e2m1x2 -> f16x2); nothing is
// dequantized to memory. Larger M goes through dequantize + the dense GEMM (ops/wo_linear.py).
//
// Reference: inference/v2/kernels/core_ops/cuda_linear (FP6-LLM W6A16 mma.sync kernel, fp6_linear.cu),
// inference/v2/kernels/cutlass_ops/mixed_gemm (W4/W8-A16 CUTLASS), csrc/fp_quantizer (selective dequantize + GEMM).
//
// CTA = 256 threads = 32 output rows x 8 K-segments of 256 elements (one 2048-element K chunk whose activations sit in
// shared memory); a thread streams its 256-element weight segment with 16-byte loads, keeps M accumulators,
// the 8 lanes of a row reduce by shuffle and one of them adds into the fp32 result (K / 2048 contributions per
// output); a second kernel converts to bf16 (+ bias).
#include "common.cuh"
#include <cuda_fp8.h>
#if __has_include(<cuda_fp4.h>)
#include <cuda_fp4.h>
#define DSB_HAVE_FP4 1
#endif
#if __has_include(<cuda_fp6.h>)
#include <cuda_fp6.h>
#define DSB_HAVE_FP6 1
#endif

namespace {

// 4 consecutive bf16 of the staged activations -> float4
__device__ __forceinline__ float4 ld_x4(const __nv_bfloat16* p)
{
    const uint2 raw = *reinterpret_cast<const uint2*>(p);
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&raw.y));
    return make_float4(a.x, a.y, b.x, b.y);
}

constexpr int WO_ROWS = 32, WO_SEGS = 8, WO_SEG = 256, WO_KCHUNK = WO_SEGS * WO_SEG;

__device__ __forceinline__ void cvt_e4m3x4(uint32_t packed, float (&f)[4])
{
    const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(packed & 0xFFFF), __NV_E4M3);
    const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(packed >> 16), __NV_E4M3);
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&lo));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    f[0] = a.x, f[1] = a.y, f[2] = b.x, f[3] = b.y;
}

__device__ __forceinline__ float e2m1_to_float(uint32_t code)
{
    // sign | 2-bit exponent | 1-bit mantissa: {0, .5, 1, 1.5, 2, 3, 4, 6}
    const uint32_t mag = code & 7u;
    const float v = (mag < 2u) ? 0.5f * (float)mag : (float)(2u | (mag & 1u)) * (float)(1u << ((mag >> 1) - 1u)) * 0.5f;
    return (code & 8u) ? -v : v;
}

// BITS: 8 (e4m3) or 4 (e2m1 nibbles, low nibble = even element)
template <int M, int BITS>
__global__ void __launch_bounds__(256) wo_gemv_kernel(const __nv_bfloat16* __restrict__ x, const uint8_t* __restrict__ wq,
                                                      const float* __restrict__ scales, float* __restrict__ y32, int N,
                                                      int K, int group, int64_t x_stride)
{
    // every 256-element segment is shifted by 8 elements: the 8 lanes of a row read 8 different segments at the same
    // offset, which would otherwise fall into the same banks
    constexpr int XROW = WO_KCHUNK + WO_SEGS * 8;
    __shared__ __align__(16) __nv_bfloat16 s_x[M][XROW];  // <= 33 KB
    const int k0 = blockIdx.y * WO_KCHUNK;
    const int kc = min(WO_KCHUNK, K - k0);  // multiple of 256
    for (int i = threadIdx.x; i < M * (WO_KCHUNK / 8); i += 256) {
        const int m = i / (WO_KCHUNK / 8), k = (i % (WO_KCHUNK / 8)) * 8;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (k < kc) v = *reinterpret_cast<const uint4*>(x + (int64_t)m * x_stride + k0 + k);
        *reinterpret_cast<uint4*>(&s_x[m][k + (k / WO_SEG) * 8]) = v;
    }
    __syncthreads();
    const int row = blockIdx.x * WO_ROWS + (threadIdx.x >> 3), seg = threadIdx.x & 7;
    float acc[M];
#pragma unroll
    for (int m = 0; m < M; m++) acc[m] = 0.f;
    if (row < N && seg * WO_SEG < kc) {
        const int64_t e0 = (int64_t)row * K + k0 + seg * WO_SEG;  // first element of this thread's segment
        const uint8_t* wp = wq + (BITS == 8 ? e0 : e0 / 2);
        constexpr int ELEMS_PER_LOAD = BITS == 8 ? 16 : 32;
        constexpr int LOADS = WO_SEG / ELEMS_PER_LOAD;
        uint4 wv[LOADS];
#pragma unroll
        for (int i = 0; i < LOADS; i++) wv[i] = __ldg(reinterpret_cast<const uint4*>(wp) + i);  // all loads in flight
#pragma unroll
        for (int i = 0; i < LOADS; i++) {
            const uint32_t words[4] = {wv[i].x, wv[i].y, wv[i].z, wv[i].w};
            const int kl = seg * (WO_SEG + 8) + i * ELEMS_PER_LOAD;  // (padded) offset inside the chunk
            float part[M];
#pragma unroll
            for (int m = 0; m < M; m++) part[m] = 0.f;
            float sc = scales[(e0 + i * ELEMS_PER_LOAD) / group];
            if constexpr (BITS == 8) {
#pragma unroll
                for (int w = 0; w < 4; w++) {
                    float f[4];
                    cvt_e4m3x4(words[w], f);
#pragma unroll
                    for (int m = 0; m < M; m++) {
                        const float4 xv = ld_x4(&s_x[m][kl + w * 4]);
                        part[m] = fmaf(f[0], xv.x, part[m]);
                        part[m] = fmaf(f[1], xv.y, part[m]);
                        part[m] = fmaf(f[2], xv.z, part[m]);
                        part[m] = fmaf(f[3], xv.w, part[m]);
                    }
                }
#pragma unroll
                for (int m = 0; m < M; m++) acc[m] = fmaf(sc, part[m], acc[m]);
            } else {
                // 32 elements; a group of >= 16 elements may end in the middle: apply the scale per 16
#pragma unroll
                for (int hsel = 0; hsel < 2; hsel++) {
                    sc = scales[(e0 + i * ELEMS_PER_LOAD + hsel * 16) / group];
#pragma unroll
                    for (int m = 0; m < M; m++) part[m] = 0.f;
#pragma unroll
                    for (int w = 0; w < 2; w++) {
                        const uint32_t word = words[hsel * 2 + w];
                        float f[8];
#pragma unroll
                        for (int n = 0; n < 8; n++) f[n] = e2m1_to_float((word >> (4 * n)) & 0xFu);
#pragma unroll
                        for (int m = 0; m < M; m++) {
                            const float4 xa = ld_x4(&s_x[m][kl + hsel * 16 + w * 8]);
                            const float4 xb = ld_x4(&s_x[m][kl + hsel * 16 + w * 8 + 4]);
                            part[m] = fmaf(f[0], xa.x, part[m]);
                            part[m] = fmaf(f[1], xa.y, part[m]);
                            part[m] = fmaf(f[2], xa.z, part[m]);
                            part[m] = fmaf(f[3], xa.w, part[m]);
                            part[m] = fmaf(f[4], xb.x, part[m]);
                            part[m] = fmaf(f[5], xb.y, part[m]);
                            part[m] = fmaf(f[6], xb.z, part[m]);
                            part[m] = fmaf(f[7], xb.w, part[m]);
                        }
                    }
#pragma unroll
                    for (int m = 0; m < M; m++) acc[m] = fmaf(sc, part[m], acc[m]);
                }
            }
        }
    }
    // the 8 lanes of a row
#pragma unroll
    for (int m = 0; m < M; m++) {
        float v = acc[m];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (seg == 0 && row < N) atomicAdd(&y32[(int64_t)m * N + row], v);
    }
}

__global__ void wo_finish_kernel(const float* __restrict__ y32, const __nv_bfloat16* __restrict__ bias,
                                 __nv_bfloat16* __restrict__ y, int M, int N, int64_t y_stride)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (int64_t)M * N) return;
    const int m = (int)(i / N), n = (int)(i % N);
    float v = y32[i];
    if (bias != nullptr) v += __bfloat162float(bias[n]);
    y[(int64_t)m * y_stride + n] = __float2bfloat16_rn(v);
}

}  // namespace

extern "C" {

// x [M, K] bf16 (row stride x_stride), wq: N*K e4m3 bytes (bits 8) or N*K/2 bytes of e2m1 nibble pairs (bits 4),
// scales fp32 [N*K / group] (group a multiple of 16 that divides K), y [M, N] bf16 (row stride y_stride),
// y32: fp32 scratch of M*N floats. M <= 8, K a multiple of 256.
int dsb200_wo_gemv(const void* x, const void* wq, const float* scales, const void* bias, void* y, float* y32, int M,
                   int N, int K, int group, int bits, int64_t x_stride, int64_t y_stride, cudaStream_t stream)
{
    if (M <= 0 || N <= 0) return 0;
    if (M > 8 || K % WO_SEG != 0 || group % 16 != 0 || K % group != 0 || (bits != 8 && bits != 4)) {
        dsb200_set_error("wo_gemv: needs M <= 8, K % 256 == 0, group % 16 == 0 dividing K, bits in {4, 8}");
        return -1;
    }
    DSB_CHECK(cudaMemsetAsync(y32, 0, (size_t)M * N * sizeof(float), stream));
    dim3 grid((N + WO_ROWS - 1) / WO_ROWS, (K + WO_KCHUNK - 1) / WO_KCHUNK);
    const int mm = M <= 1 ? 1 : (M <= 2 ? 2 : (M <= 4 ? 4 : 8));
    const __nv_bfloat16* xp = (const __nv_bfloat16*)x;
    const uint8_t* wp = (const uint8_t*)wq;
    // rows M..mm-1 of the template read x rows that do not exist: clamp by launching the exact-M kernel when possible
#define WO_LAUNCH(MM, BB) wo_gemv_kernel<MM, BB><<<grid, 256, 0, stream>>>(xp, wp, scales, y32, N, K, group, x_stride)
#define WO_BITS(MM)                       \
    do {                                  \
        if (bits == 8) WO_LAUNCH(MM, 8);  \
        else WO_LAUNCH(MM, 4);            \
    } while (0)
    (void)mm;
    switch (M) {
        case 1: WO_BITS(1); break;
        case 2: WO_BITS(2); break;
        case 3: WO_BITS(3); break;
        case 4: WO_BITS(4); break;
        case 5: WO_BITS(5); break;
        case 6: WO_BITS(6); break;
        case 7: WO_BITS(7); break;
        default: WO_BITS(8); break;
    }
#undef WO_BITS
#undef WO_LAUNCH
    DSB_CHECK_LAUNCH("wo_gemv");
    const int64_t total = (int64_t)M * N;
    wo_finish_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(y32, (const __nv_bfloat16*)bias,
                                                                          (__nv_bfloat16*)y, M, N, y_stride);
    DSB_CHECK_LAUNCH("wo_finish");
    return 0;
}

}  // extern "C"

// =========================================================================================================
// tcgen05 version: Y[T, N] = X[T, K] * dequant(Wq[N, K])^T for T <= 64 tokens, computed as D = W * X^T so that the
// (large) weight matrix is the M operand of the MMA and the few tokens are its N extent (16 / 32 / 64).
//
//   warp 0      TMA: raw quantized weight tile (128 rows x 64 elements: 8 KB FP8 / 6 KB FP6 / 4 KB FP4, unswizzled) and the
//               activation tile (TN tokens x 64, bf16, 128-byte swizzle) of every k-block into a 4-stage ring
//   warps 6-13  expand: raw codes -> f16 with ONE hardware cvt per two elements (cvt.f16x2.e4m3x2 / .e3m2x2 / .e2m1x2, exact),
//               written straight into the 128-byte-swizzled K-major layout the tensor core reads; no scale here
//   warp 1      one lane issues 4 x tcgen05.mma (A = f16 weights, B = activations converted to f16, M128 x N{16,32,64} x K16) per
//               k-block; each scale group (`group` elements of K) accumulates into its own TMEM buffer (two alternate)
//   warps 2-5   per group: tcgen05.ld the group's partial sums, multiply by the fp32 group scale of the row, add into
//               fp32 registers; at the end bf16 store (+ bias), or fp32 atomics when split-K CTAs share an output
//
// The weights cross HBM once in quantized form and cost ~0.7 SIMT instructions per element (the scale is applied to
// the per-group partial sums in fp32, which is also more accurate than scaling the weights to bf16 first).
// Grid = (N / 128) x split-K so that all SMs stream.
// =========================================================================================================
#include "gemm_common.cuh"

namespace {

// ~48 KB of raw weights in flight per CTA (two CTAs per SM): HBM latency x bandwidth needs ~90 KB per SM; ncu showed the
// 4-stage version waiting on the ring (37 % of DRAM peak, long-scoreboard stalls)
constexpr int WT_RING_BYTES = 48 * 1024;
constexpr int WT_MAX_STAGES = 12;
constexpr int WT_ABUFS = 2;  // expanded weight tiles in flight between the expand warps and the MMA
constexpr int WT_BM = 128, WT_BK = 64;
constexpr int WT_THREADS = 64 + 128 + 256;  // TMA warp, MMA warp, 4 accumulate / epilogue warps, 8 expand warps

struct WtBars {
    uint64_t raw_full[WT_MAX_STAGES], raw_empty[WT_MAX_STAGES];
    uint64_t a_full[WT_ABUFS], a_empty[WT_ABUFS];
    uint64_t acc_full[2], acc_empty[2];
    uint32_t tmem_base;
};

// kind::f16 with A = B = f16, D = f32, both K-major (mixing an f16 with a bf16 operand is an illegal instruction)
__host__ __device__ constexpr uint32_t make_idesc_f16(int m, int n)
{
    return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// activations bf16 -> f16 (saturating at +-65504), so that both MMA operands are f16
__global__ void __launch_bounds__(256) wo_x_to_f16_kernel(const __nv_bfloat16* __restrict__ x, __half* __restrict__ x16, int T,
                                                          int K, int64_t x_stride)
{
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i >= (int64_t)T * K) return;
    const int t = (int)(i / K), k = (int)(i % K);
    const uint4 raw = *reinterpret_cast<const uint4*>(x + (int64_t)t * x_stride + k);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
    uint4 out;
    __half2* o = reinterpret_cast<__half2*>(&out);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float2 f = __bfloat1622float2(h[j]);
        f.x = fminf(fmaxf(f.x, -65504.f), 65504.f);
        f.y = fminf(fmaxf(f.y, -65504.f), 65504.f);
        o[j] = __floats2half2_rn(f.x, f.y);
    }
    *reinterpret_cast<uint4*>(x16 + i) = out;
}

template <int BITS, int TN>
__global__ void __launch_bounds__(WT_THREADS, 2)
wo_gemm_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_x,
                  const float* __restrict__ scales, float* __restrict__ y32, const __nv_bfloat16* __restrict__ bias,
                  __nv_bfloat16* __restrict__ y, int64_t y_stride, int T, int N, int K, int group, int kb_per_split)
{
    constexpr int RAW_BYTES = WT_BM * WT_BK * BITS / 8;  // 8 KB / 6 KB / 4 KB
    constexpr int WT_STAGES = WT_RING_BYTES / RAW_BYTES;  // 6 / 8 / 12
    constexpr int X_BYTES_MAX = (TN * WT_BK * 2 + 1023) / 1024 * 1024;  // 2 / 4 / 8 KB
    constexpr int A_BYTES_T = WT_BM * WT_BK * 2;         // 16 KB
    extern __shared__ uint8_t smem_raw[];
    uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* a_smem = base;                                   // WT_ABUFS x 16 KB (swizzled f16 weight tiles)
    uint8_t* x_smem = a_smem + WT_ABUFS * A_BYTES_T;          // WT_STAGES x X_BYTES_MAX
    uint8_t* r_smem = x_smem + WT_STAGES * X_BYTES_MAX;       // WT_STAGES x RAW_BYTES
    WtBars* bars = reinterpret_cast<WtBars*>(r_smem + WT_STAGES * RAW_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * WT_BM;
    const int kb0 = blockIdx.y * kb_per_split;
    const int kb_total = K / WT_BK;
    const int nkb = min(kb_per_split, kb_total - kb0);
    if (nkb <= 0) return;
    const int kbg = group / WT_BK;  // k-blocks per scale group (group is a multiple of 64)
    // groups touched by this CTA: absolute group index of k-block kb is (kb0 + kb) / kbg
    const int g_first = kb0 / kbg, g_last = (kb0 + nkb - 1) / kbg;

    if (threadIdx.x == 0) {
        for (int s = 0; s < WT_STAGES; s++) {
            mbar_init(&bars->raw_full[s], 1);
            mbar_init(&bars->raw_empty[s], 1);
        }
        for (int s = 0; s < WT_ABUFS; s++) {
            mbar_init(&bars->a_full[s], 8);
            mbar_init(&bars->a_empty[s], 1);
        }
        for (int s = 0; s < 2; s++) {
            mbar_init(&bars->acc_full[s], 1);
            mbar_init(&bars->acc_empty[s], 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_x) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "n"(128)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    const uint32_t x_bytes = (uint32_t)TN * WT_BK * 2;

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = 0; kb < nkb; kb++) {
                mbar_wait(&bars->raw_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->raw_full[stage], RAW_BYTES + x_bytes);
                // weights: payload viewed as [N rows, K * BITS / 8 bytes]; box = [64 * BITS / 8 bytes, 128 rows]
                tma_load_2d(r_smem + stage * RAW_BYTES, &map_w, &bars->raw_full[stage], (kb0 + kb) * (WT_BK * BITS / 8), n0);
                tma_load_2d(x_smem + stage * X_BYTES_MAX, &map_x, &bars->raw_full[stage], (kb0 + kb) * WT_BK, 0);
                if (++stage == WT_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_f16(WT_BM, TN);
            int stage = 0;
            uint32_t phase = 0;
            int cur_g = -1, gi = -1;
            for (int kb = 0; kb < nkb; kb++) {
                const int a = kb % WT_ABUFS;
                const int g = (kb0 + kb) / kbg;
                bool first_of_group = false;
                if (g != cur_g) {  // a new scale group starts: switch accumulator buffer
                    if (cur_g >= 0) umma_commit(&bars->acc_full[gi & 1]);
                    cur_g = g;
                    gi++;
                    mbar_wait(&bars->acc_empty[gi & 1], (uint32_t)(((gi >> 1) & 1) ^ 1));
                    tc_fence_after();
                    first_of_group = true;
                }
                mbar_wait(&bars->raw_full[stage], phase);                        // activation tile landed
                mbar_wait(&bars->a_full[a], (uint32_t)((kb / WT_ABUFS) & 1));    // weight tile expanded
                tc_fence_after();
                const uint64_t adesc = make_smem_desc(smem_u32(a_smem + a * A_BYTES_T));
                const uint64_t bdesc = make_smem_desc(smem_u32(x_smem + stage * X_BYTES_MAX));
                const uint32_t d_addr = tmem_base + (uint32_t)((gi & 1) * 64);
#pragma unroll
                for (int k = 0; k < WT_BK / UMMA_K; k++)
                    umma_bf16(d_addr, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                              (!first_of_group || k > 0) ? 1u : 0u);
                umma_commit(&bars->raw_empty[stage]);
                umma_commit(&bars->a_empty[a]);
                if (++stage == WT_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
            umma_commit(&bars->acc_full[gi & 1]);
        }
    } else if (warp >= 6) {
        // ---------------- expand raw codes to f16 (8 warps: two (row, 16-element chunk) units per thread) ----------------
        const int tq = threadIdx.x - 192;  // 0..255
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < nkb; kb++) {
            const int a = kb % WT_ABUFS;
            mbar_wait(&bars->raw_full[stage], phase);
            mbar_wait(&bars->a_empty[a], (uint32_t)(((kb / WT_ABUFS) & 1) ^ 1));
            const uint8_t* raw = r_smem + stage * RAW_BYTES;
            uint8_t* dst = a_smem + a * A_BYTES_T;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                const int row = i * 64 + (tq >> 2), c4 = tq & 3;  // 16 elements: logical 16-byte chunks 2*c4, 2*c4+1
                uint32_t out[8];
                if constexpr (BITS == 8) {
                    const uint4 rv = *reinterpret_cast<const uint4*>(raw + row * 64 + c4 * 16);
                    const uint32_t words[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
                    for (int w = 0; w < 4; w++) {
                        const __half2_raw lo = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(words[w] & 0xFFFF), __NV_E4M3);
                        const __half2_raw hi = __nv_cvt_fp8x2_to_halfraw2((__nv_fp8x2_storage_t)(words[w] >> 16), __NV_E4M3);
                        out[2 * w] = *reinterpret_cast<const uint32_t*>(&lo);
                        out[2 * w + 1] = *reinterpret_cast<const uint32_t*>(&hi);
                    }
                } else if constexpr (BITS == 6) {
#ifdef DSB_HAVE_FP6
                    // 16 e3m2 codes = 96 bits (code i at bits [6i, 6i+6) of the row's little-endian bit stream): each
                    // 12-bit pair is spread to two bytes and converted by one cvt.rn.f16x2.e3m2x2
                    const uint32_t* rp = reinterpret_cast<const uint32_t*>(raw + row * 48 + c4 * 12);
                    const uint32_t w0 = rp[0], w1 = rp[1], w2 = rp[2];
                    const uint32_t pr[8] = {w0,
                                            w0 >> 12,
                                            __funnelshift_r(w0, w1, 24),
                                            w1 >> 4,
                                            w1 >> 16,
                                            __funnelshift_r(w1, w2, 28),
                                            w2 >> 8,
                                            w2 >> 20};
#pragma unroll
                    for (int p = 0; p < 8; p++) {
                        const uint32_t v = pr[p];
                        const __half2_raw h2 =
                            __nv_cvt_fp6x2_to_halfraw2((__nv_fp6x2_storage_t)((v & 0x3F) | ((v & 0xFC0) << 2)), __NV_E3M2);
                        out[p] = *reinterpret_cast<const uint32_t*>(&h2);
                    }
#endif
                } else {
#ifdef DSB_HAVE_FP4
                    const uint2 rv = *reinterpret_cast<const uint2*>(raw + row * 32 + c4 * 8);
                    const uint32_t words[2] = {rv.x, rv.y};
#pragma unroll
                    for (int w = 0; w < 2; w++)
#pragma unroll
                        for (int p = 0; p < 4; p++) {
                            const __half2_raw h2 =
                                __nv_cvt_fp4x2_to_halfraw2((__nv_fp4x2_storage_t)((words[w] >> (8 * p)) & 0xFF), __NV_E2M1);
                            out[w * 4 + p] = *reinterpret_cast<const uint32_t*>(&h2);
                        }
#endif
                }
                // 128-byte swizzle of the K-major tile: 16-byte chunk c of row r lives at chunk (c ^ (r & 7))
                const int sw = row & 7;
                *reinterpret_cast<uint4*>(dst + row * 128 + (((2 * c4) ^ sw) << 4)) = make_uint4(out[0], out[1], out[2], out[3]);
                *reinterpret_cast<uint4*>(dst + row * 128 + (((2 * c4 + 1) ^ sw) << 4)) = make_uint4(out[4], out[5], out[6], out[7]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->a_full[a]);
            if (++stage == WT_STAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else {
        // ---------------- per-group accumulate + epilogue (warps 2-5 own the four TMEM lane groups) ----------------
        const int lane_group = warp & 3;
        const int n = n0 + lane_group * 32 + lane;
        float acc[TN];
#pragma unroll
        for (int j = 0; j < TN; j++) acc[j] = 0.f;
        int gi = 0;
        for (int g = g_first; g <= g_last; g++, gi++) {
            const float sc = (n < N) ? __ldg(scales + (int64_t)n * (K / group) + g) : 0.f;
            mbar_wait(&bars->acc_full[gi & 1], (uint32_t)((gi >> 1) & 1));
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)((gi & 1) * 64);
#pragma unroll
            for (int c = 0; c < TN; c += 16) {
                {
                    uint32_t r[16];
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
                        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]),
                          "=r"(r[15])
                        : "r"(taddr + c)
                        : "memory");
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                    for (int j = 0; j < 16; j++) acc[c + j] = fmaf(sc, __uint_as_float(r[j]), acc[c + j]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->acc_empty[gi & 1]);
        }
        if (n < N) {
            const bool direct = gridDim.y == 1;  // no split-K: write bf16 (+ bias) straight to y
            const float bv = (direct && bias != nullptr) ? __bfloat162float(bias[n]) : 0.f;
#pragma unroll
            for (int j = 0; j < TN; j++) {
                if (j < T) {
                    if (direct)
                        y[(int64_t)j * y_stride + n] = __float2bfloat16_rn(acc[j] + bv);
                    else
                        atomicAdd(&y32[(int64_t)j * N + n], acc[j]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(128) : "memory");
    }
}

// plain 2-D tensor map over bytes (no swizzle): [rows, row_bytes], box = [box_bytes, box_rows]
int make_map_bytes(CUtensorMap* map, const void* ptr, int64_t rows, int64_t row_bytes, int box_bytes, int box_rows)
{
    EncodeTiledFn enc = get_encode();
    if (enc == nullptr) {
        dsb200_set_error("cuTensorMapEncodeTiled entry point unavailable");
        return -1;
    }
    cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
    cuuint32_t box[2] = {(cuuint32_t)box_bytes, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        dsb200_set_error("wo_gemm: cuTensorMapEncodeTiled failed for the weight payload (16-byte aligned rows needed)");
        return -2;
    }
    return 0;
}

}  // namespace

extern "C" {

// As dsb200_wo_gemv, on the tensor cores, for T <= 64 tokens. K a multiple of 64; the scale group a multiple of 64 that
// divides K (scales laid out [N, K / group]).
// x16: scratch of T * K halves (the activations as f16); y32: T * N floats (used when split-K is taken).
int dsb200_wo_gemm_tc(const void* x, const void* wq, const float* scales, const void* bias, void* y, float* y32, void* x16,
                      int T, int N, int K, int group, int bits, int64_t x_stride, int64_t y_stride, cudaStream_t stream)
{
    if (T <= 0 || N <= 0) return 0;
    if (T > 64 || K % WT_BK != 0 || group % WT_BK != 0 || K % group != 0 || (bits != 8 && bits != 6 && bits != 4) ||
        (x_stride % 8) || ((uintptr_t)x & 15) || ((uintptr_t)wq & 15)) {
        dsb200_set_error("wo_gemm_tc: needs T <= 64, K % 64 == 0, group % 64 == 0 dividing K, bits in {4, 6, 8}, aligned operands");
        return -1;
    }
#ifndef DSB_HAVE_FP6
    if (bits == 6) {
        dsb200_set_error("wo_gemm_tc: built without cuda_fp6.h");
        return -1;
    }
#endif
#ifndef DSB_HAVE_FP4
    if (bits == 4) {
        dsb200_set_error("wo_gemm_tc: built without cuda_fp4.h");
        return -1;
    }
#endif
    const int TN = T <= 16 ? 16 : (T <= 32 ? 32 : 64);
    CUtensorMap mw, mx;
    if (make_map_bytes(&mw, wq, N, (int64_t)K * bits / 8, WT_BK * bits / 8, WT_BM) != 0) return -2;
    wo_x_to_f16_kernel<<<(unsigned)(((int64_t)T * K / 8 + 255) / 256), 256, 0, stream>>>((const __nv_bfloat16*)x, (__half*)x16,
                                                                                        T, K, x_stride);
    DSB_CHECK_LAUNCH("wo_x_to_f16");
    if (make_map(&mx, x16, T, K, K, TN) != 0) return -2;  // same element size / swizzle as a bf16 map
    const int n_blocks = (N + WT_BM - 1) / WT_BM, kb_total = K / WT_BK;
    int splits = (2 * DSB_NUM_SMS) / n_blocks;  // aim at ~2 CTAs per SM in total; none when N alone fills the GPU
    if (splits > kb_total / 4) splits = kb_total / 4 > 0 ? kb_total / 4 : 1;
    if (splits < 1) splits = 1;
    const int kb_per_split = (kb_total + splits - 1) / splits;
    splits = (kb_total + kb_per_split - 1) / kb_per_split;
    if (splits > 1) DSB_CHECK(cudaMemsetAsync(y32, 0, (size_t)T * N * sizeof(float), stream));
    dim3 grid(n_blocks, splits);
    const int raw_bytes = WT_BM * WT_BK * bits / 8;
    const int x_stage = (TN * WT_BK * 2 + 1023) / 1024 * 1024;
    const int n_stages = WT_RING_BYTES / raw_bytes;
    const int smem = WT_ABUFS * WT_BM * WT_BK * 2 + n_stages * x_stage + n_stages * raw_bytes + 1024 + 512;
    int dev_ = 0;
    cudaGetDevice(&dev_);
#define WT_LAUNCH_(BB, TT)                                                                                             \
    do {                                                                                                               \
        static bool attr_set[64] = {};                                                                                 \
        if (!attr_set[dev_ & 63]) {                                                                                    \
            DSB_CHECK(cudaFuncSetAttribute(wo_gemm_tc_kernel<BB, TT>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                           smem));                                                                     \
            attr_set[dev_ & 63] = true;                                                                                \
        }                                                                                                              \
        wo_gemm_tc_kernel<BB, TT><<<grid, WT_THREADS, smem, stream>>>(mw, mx, scales, y32, (const __nv_bfloat16*)bias, \
                                                                      (__nv_bfloat16*)y, y_stride, T, N, K, group,      \
                                                                      kb_per_split);                                   \
    } while (0)
#define WT_BITS(TT)            \
    do {                       \
        if (bits == 8) {       \
            constexpr int BB = 8; \
            WT_LAUNCH_(BB, TT); \
        } else if (bits == 6) { \
            constexpr int BB = 6; \
            WT_LAUNCH_(BB, TT); \
        } else {               \
            constexpr int BB = 4; \
            WT_LAUNCH_(BB, TT); \
        }                      \
    } while (0)
    if (TN == 16)
        WT_BITS(16);
    else if (TN == 32)
        WT_BITS(32);
    else
        WT_BITS(64);
#undef WT_BITS
#undef WT_LAUNCH_
    DSB_CHECK_LAUNCH("wo_gemm_tc");
    if (splits > 1) {
        const int64_t total = (int64_t)T * N;
        wo_finish_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(y32, (const __nv_bfloat16*)bias,
                                                                              (__nv_bfloat16*)y, T, N, y_stride);
        DSB_CHECK_LAUNCH("wo_finish");
    }
    return 0;
}

}  // extern "C"
