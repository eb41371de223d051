// Flash-attention backward on the 5th-generation tensor cores (bf16, head_dim 128, causal or full, GQA, Sq == Skv).
//
//   delta[b, h, i] = sum_d dO[i, d] * O[i, d]                                  (preprocess, SIMT)
//   dQ kernel   one CTA per (batch, head, 128-row query block), loops over the key blocks:
//               S = Q K^T, dP = dO V^T  ->  P = exp2(S*scale - LSE), dS = P o (dP - delta) * scale  ->  dQ += dS K
//   dK/dV kernel one CTA per (batch, KV head, 128-key block), loops over the query heads of the group and their query
//               blocks, everything transposed so that the softmax-side tiles are born in the operand orientation:
//               S^T = K Q^T, dP^T = V dO^T  ->  P^T, dS^T  ->  dV += P^T dO, dK += dS^T Q
// Both recompute S / dP (7 tile products instead of 5) -- in exchange there are no atomics, no fp32 dQ scratch and the
// result is deterministic. P / dS never leave TMEM: the softmax warps overwrite the fp32 S / dP columns with bf16 and the
// next MMA reads them as its A operand (TS form); Q / dO / K / V tiles serve both as K-major and, through a second
// descriptor on the same bytes, as MN-major operands.
//
// Reference: the reference trains through the framework's flash-attention backward
// (sequence/fpdt_layer.py:238-269 _flash_attn_backward; HF attention functions); its own softmax backward kernels
// (csrc/transformer/softmax_kernels.cu softmax_backward_kernel*) need the materialised [S, S] probabilities.
#include "attention_common.cuh"

namespace {

constexpr int FB_THREADS = 192;  // TMA warp, MMA warp, 4 softmax / epilogue warps

struct FbParams {
    const float* lse;    // [B, H, S] natural log
    const float* delta;  // [B, H, S]
    __nv_bfloat16* dq;   // [B, S, H, 128]
    __nv_bfloat16* dk;   // [B, S, Hkv, 128]
    __nv_bfloat16* dv;
    int B, S, H, Hkv;
    int64_t dq_row_stride, dq_batch_stride, dk_row_stride, dk_batch_stride, dv_row_stride, dv_batch_stride;
    float scale, scale_log2;
    int causal;
    int nblk;  // S / 128
};

__device__ __forceinline__ void named_bar_sync(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// ------------------------------------------------------------------------------------------------------------
// preprocess: delta = rowsum(dO * O); one warp per (b, i, h) row of 128 elements
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fa_bwd_delta_kernel(const __nv_bfloat16* __restrict__ o,
                                                           const __nv_bfloat16* __restrict__ dout,
                                                           const float* __restrict__ dlse,
                                                           float* __restrict__ delta, int B, int S, int H,
                                                           int64_t o_row, int64_t o_batch, int64_t do_row,
                                                           int64_t do_batch)
{
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= (int64_t)B * S * H) return;
    const int h = (int)(w % H);
    const int64_t bi = w / H;
    const int i = (int)(bi % S), b = (int)(bi / S);
    const uint2 a = *reinterpret_cast<const uint2*>(o + b * o_batch + (int64_t)i * o_row + h * 128 + lane * 4);
    const uint2 g = *reinterpret_cast<const uint2*>(dout + b * do_batch + (int64_t)i * do_row + h * 128 + lane * 4);
    const float2 a0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&a.x));
    const float2 a1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&a.y));
    const float2 g0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&g.x));
    const float2 g1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&g.y));
    float v = a0.x * g0.x + a0.y * g0.y + a1.x * g1.x + a1.y * g1.y;
    v = dsb::warp_sum(v);
    // a gradient flowing into the log-sum-exp output (chunked / ring attention merges partial results through it) enters
    // dS = P o (dP - delta) as an extra P o dLSE term, i.e. as a shift of delta
    if (lane == 0) {
        const int64_t idx = ((int64_t)b * H + h) * S + i;
        delta[idx] = v - (dlse != nullptr ? dlse[idx] : 0.f);
    }
}

// ------------------------------------------------------------------------------------------------------------
// dQ kernel
// ------------------------------------------------------------------------------------------------------------
struct DqBars {
    uint64_t q_full;
    uint64_t kv_full[2], kv_empty[2];
    uint64_t s_full, p_full, d_full;
    uint32_t tmem_base;
};
constexpr int DQ_SMEM_BYTES = 6 * FA_TILE_BYTES + 1024 + 256;  // Q, dO, 2 x (K, V)

__global__ void __launch_bounds__(FB_THREADS, 1)
fa_bwd_dq_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_do,
                 const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v, const FbParams prm)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_smem = tiles;
    uint8_t* do_smem = tiles + FA_TILE_BYTES;
    uint8_t* kv_smem = tiles + 2 * FA_TILE_BYTES;  // stage s: K at +0, V at +FA_TILE_BYTES
    DqBars* bars = reinterpret_cast<DqBars*>(tiles + 6 * FA_TILE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // heavy (late) query blocks first
    const int bh = prm.B * prm.H;
    const int i = prm.nblk - 1 - (int)(blockIdx.x / bh);
    const int rem = (int)(blockIdx.x % bh);
    const int b = rem / prm.H, h = rem % prm.H;
    const int hk = h / (prm.H / prm.Hkv);
    const int nb = prm.causal ? i + 1 : prm.nblk;  // key blocks 0..nb-1

    if (threadIdx.x == 0) {
        mbar_init(&bars->q_full, 1);
        for (int s = 0; s < 2; s++) {
            mbar_init(&bars->kv_full[s], 1);
            mbar_init(&bars->kv_empty[s], 1);
        }
        mbar_init(&bars->s_full, 1);
        mbar_init(&bars->p_full, 4);
        mbar_init(&bars->d_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "n"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    constexpr uint32_t COL_S = 0, COL_DP = 128, COL_DQ = 256;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(&bars->q_full, 2 * FA_TILE_BYTES);
#pragma unroll
            for (int hf = 0; hf < 2; hf++) {
                tma_load_3d(q_smem + hf * FA_HALF_BYTES, &map_q, &bars->q_full, h * FA_D + hf * 64, i * FA_BM, b);
                tma_load_3d(do_smem + hf * FA_HALF_BYTES, &map_do, &bars->q_full, h * FA_D + hf * 64, i * FA_BM, b);
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < nb; j++) {
                mbar_wait(&bars->kv_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->kv_full[stage], 2 * FA_TILE_BYTES);
                uint8_t* kd = kv_smem + stage * 2 * FA_TILE_BYTES;
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    tma_load_3d(kd + hf * FA_HALF_BYTES, &map_k, &bars->kv_full[stage], hk * FA_D + hf * 64, j * FA_BN, b);
                    tma_load_3d(kd + FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_v, &bars->kv_full[stage], hk * FA_D + hf * 64,
                                j * FA_BN, b);
                }
                if (++stage == 2) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_kk = make_idesc_ex(128, 128, 0, 0);  // both operands K-major
            constexpr uint32_t idesc_mn = make_idesc_ex(128, 128, 0, 1);  // B operand MN-major
            const uint32_t qa = smem_u32(q_smem), da = smem_u32(do_smem);
            auto issue_s_dp = [&](int stage) {
                const uint32_t ka = smem_u32(kv_smem + stage * 2 * FA_TILE_BYTES), va = ka + FA_TILE_BYTES;
#pragma unroll
                for (int k = 0; k < FA_D / UMMA_K; k++) {
                    const uint32_t off = (uint32_t)(k >> 2) * FA_HALF_BYTES + (uint32_t)(k & 3) * 32;
                    umma_bf16(tmem_base + COL_S, make_smem_desc(qa + off), make_smem_desc(ka + off), idesc_kk, k > 0 ? 1u : 0u);
                }
#pragma unroll
                for (int k = 0; k < FA_D / UMMA_K; k++) {
                    const uint32_t off = (uint32_t)(k >> 2) * FA_HALF_BYTES + (uint32_t)(k & 3) * 32;
                    umma_bf16(tmem_base + COL_DP, make_smem_desc(da + off), make_smem_desc(va + off), idesc_kk, k > 0 ? 1u : 0u);
                }
                umma_commit(&bars->s_full);
            };
            mbar_wait(&bars->q_full, 0);
            mbar_wait(&bars->kv_full[0], 0);
            tc_fence_after();
            issue_s_dp(0);
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < nb; j++) {
                mbar_wait(&bars->p_full, (uint32_t)(j & 1));
                tc_fence_after();
                // dQ += dS K_j : A = dS (bf16 in TMEM over the dP columns), B = K_j as [N = d, K = keys] (MN-major)
                const uint32_t ka = smem_u32(kv_smem + stage * 2 * FA_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < FA_BN / UMMA_K; k++)
                    umma_bf16_ts(tmem_base + COL_DQ, tmem_base + COL_DP + k * 8, make_smem_desc_mn(ka + k * 2048, FA_HALF_BYTES),
                                 idesc_mn, (j > 0 || k > 0) ? 1u : 0u);
                umma_commit(&bars->kv_empty[stage]);
                if (++stage == 2) {
                    stage = 0;
                    phase ^= 1;
                }
                if (j + 1 < nb) {
                    mbar_wait(&bars->kv_full[stage], phase);
                    tc_fence_after();
                    issue_s_dp(stage);
                } else {
                    umma_commit(&bars->d_full);
                }
            }
        }
    } else {
        const int lane_base = (warp & 3) * 32;
        const int row = i * FA_BM + lane_base + lane;  // query index
        const uint32_t lane_addr = (uint32_t)lane_base << 16;
        const uint32_t s_addr = tmem_base + lane_addr + COL_S, dp_addr = tmem_base + lane_addr + COL_DP;
        const int64_t stat = ((int64_t)b * prm.H + h) * prm.S + row;
        const float lse2 = prm.lse[stat] * 1.4426950408889634f;
        const float dlt = prm.delta[stat];
        for (int j = 0; j < nb; j++) {
            mbar_wait(&bars->s_full, (uint32_t)(j & 1));
            tc_fence_after();
            const bool diag = prm.causal && j == i;
            const int col0 = j * FA_BN;
#pragma unroll 1
            for (int c = 0; c < 4; c++) {
                uint32_t s[32], dp[32], pk[16];
                tmem_ld_32x32(s_addr + c * 32, s);
                tmem_ld_32x32(dp_addr + c * 32, dp);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 16; e++) {
                    float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * e]), prm.scale_log2, -lse2));
                    float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * e + 1]), prm.scale_log2, -lse2));
                    if (diag) {
                        if (col0 + c * 32 + 2 * e > row) p0 = 0.f;
                        if (col0 + c * 32 + 2 * e + 1 > row) p1 = 0.f;
                    }
                    const float d0 = p0 * (__uint_as_float(dp[2 * e]) - dlt) * prm.scale;
                    const float d1 = p1 * (__uint_as_float(dp[2 * e + 1]) - dlt) * prm.scale;
                    __nv_bfloat162 hh = __floats2bfloat162_rn(d0, d1);
                    pk[e] = *reinterpret_cast<uint32_t*>(&hh);
                }
                tmem_st_32x16(dp_addr + c * 16, pk);  // dS over the (already consumed) dP columns
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->p_full);
        }
        // epilogue: dQ -> bf16
        mbar_wait(&bars->d_full, 0);
        tc_fence_after();
        __nv_bfloat16* drow = prm.dq + (int64_t)b * prm.dq_batch_stride + (int64_t)row * prm.dq_row_stride + h * FA_D;
#pragma unroll 1
        for (int c = 0; c < 4; c++) {
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + lane_addr + COL_DQ + c * 32, r);
            tmem_wait_ld();
#pragma unroll
            for (int v4 = 0; v4 < 4; v4++) {
                uint4 pk;
                __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int e = 0; e < 4; e++)
                    hh[e] = __floats2bfloat162_rn(__uint_as_float(r[v4 * 8 + 2 * e]), __uint_as_float(r[v4 * 8 + 2 * e + 1]));
                *reinterpret_cast<uint4*>(drow + c * 32 + v4 * 8) = pk;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------------------
// dK / dV kernel
// ------------------------------------------------------------------------------------------------------------
struct DkvBars {
    uint64_t kv_full;
    uint64_t qdo_full[2], qdo_empty[2];
    uint64_t s_full, p_full, d_full;
    uint32_t tmem_base;
    alignas(16) float stats[2][2][128];  // [buffer][lse2 | delta][query row of the block]
};
constexpr int DKV_SMEM_BYTES = 6 * FA_TILE_BYTES + 1024 + 512 + 2 * 2 * 128 * 4;  // K, V, 2 x (Q, dO)

__global__ void __launch_bounds__(FB_THREADS, 1)
fa_bwd_dkv_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_do,
                  const __grid_constant__ CUtensorMap map_k, const __grid_constant__ CUtensorMap map_v, const FbParams prm)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* k_smem = tiles;
    uint8_t* v_smem = tiles + FA_TILE_BYTES;
    uint8_t* qdo_smem = tiles + 2 * FA_TILE_BYTES;  // stage s: Q at +0, dO at +FA_TILE_BYTES
    DkvBars* bars = reinterpret_cast<DkvBars*>(tiles + 6 * FA_TILE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // key blocks with the most query blocks behind them (early ones under the causal mask) first
    const int bhk = prm.B * prm.Hkv;
    const int j = (int)(blockIdx.x / bhk);
    const int rem = (int)(blockIdx.x % bhk);
    const int b = rem / prm.Hkv, hk = rem % prm.Hkv;
    const int G = prm.H / prm.Hkv;
    const int i0 = prm.causal ? j : 0;
    const int per_head = prm.nblk - i0;  // query blocks per head
    const int n_it = G * per_head;

    if (threadIdx.x == 0) {
        mbar_init(&bars->kv_full, 1);
        for (int s = 0; s < 2; s++) {
            mbar_init(&bars->qdo_full[s], 1);
            mbar_init(&bars->qdo_empty[s], 1);
        }
        mbar_init(&bars->s_full, 1);
        mbar_init(&bars->p_full, 4);
        mbar_init(&bars->d_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&bars->tmem_base)),
                     "n"(512)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_base;
    constexpr uint32_t COL_ST = 0, COL_DPT = 128, COL_DV = 256, COL_DK = 384;

    if (warp == 0) {
        if (lane == 0) {
            mbar_expect_tx(&bars->kv_full, 2 * FA_TILE_BYTES);
#pragma unroll
            for (int hf = 0; hf < 2; hf++) {
                tma_load_3d(k_smem + hf * FA_HALF_BYTES, &map_k, &bars->kv_full, hk * FA_D + hf * 64, j * FA_BN, b);
                tma_load_3d(v_smem + hf * FA_HALF_BYTES, &map_v, &bars->kv_full, hk * FA_D + hf * 64, j * FA_BN, b);
            }
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < n_it; it++) {
                const int h = hk * G + it / per_head, i = i0 + it % per_head;
                mbar_wait(&bars->qdo_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->qdo_full[stage], 2 * FA_TILE_BYTES);
                uint8_t* qd = qdo_smem + stage * 2 * FA_TILE_BYTES;
#pragma unroll
                for (int hf = 0; hf < 2; hf++) {
                    tma_load_3d(qd + hf * FA_HALF_BYTES, &map_q, &bars->qdo_full[stage], h * FA_D + hf * 64, i * FA_BM, b);
                    tma_load_3d(qd + FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_do, &bars->qdo_full[stage], h * FA_D + hf * 64,
                                i * FA_BM, b);
                }
                if (++stage == 2) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc_kk = make_idesc_ex(128, 128, 0, 0);
            constexpr uint32_t idesc_mn = make_idesc_ex(128, 128, 0, 1);
            const uint32_t ka = smem_u32(k_smem), va = smem_u32(v_smem);
            auto issue_st_dpt = [&](int stage) {
                const uint32_t qa = smem_u32(qdo_smem + stage * 2 * FA_TILE_BYTES), da = qa + FA_TILE_BYTES;
#pragma unroll
                for (int k = 0; k < FA_D / UMMA_K; k++) {  // S^T = K Q^T
                    const uint32_t off = (uint32_t)(k >> 2) * FA_HALF_BYTES + (uint32_t)(k & 3) * 32;
                    umma_bf16(tmem_base + COL_ST, make_smem_desc(ka + off), make_smem_desc(qa + off), idesc_kk, k > 0 ? 1u : 0u);
                }
#pragma unroll
                for (int k = 0; k < FA_D / UMMA_K; k++) {  // dP^T = V dO^T
                    const uint32_t off = (uint32_t)(k >> 2) * FA_HALF_BYTES + (uint32_t)(k & 3) * 32;
                    umma_bf16(tmem_base + COL_DPT, make_smem_desc(va + off), make_smem_desc(da + off), idesc_kk, k > 0 ? 1u : 0u);
                }
                umma_commit(&bars->s_full);
            };
            mbar_wait(&bars->kv_full, 0);
            mbar_wait(&bars->qdo_full[0], 0);
            tc_fence_after();
            issue_st_dpt(0);
            int stage = 0;
            uint32_t phase = 0;
            for (int it = 0; it < n_it; it++) {
                mbar_wait(&bars->p_full, (uint32_t)(it & 1));
                tc_fence_after();
                const uint32_t qa = smem_u32(qdo_smem + stage * 2 * FA_TILE_BYTES), da = qa + FA_TILE_BYTES;
#pragma unroll
                for (int k = 0; k < FA_BM / UMMA_K; k++) {
                    // dV += P^T dO : A = P^T (TMEM, K = query rows), B = dO as [N = d, K = query rows] (MN-major)
                    umma_bf16_ts(tmem_base + COL_DV, tmem_base + COL_ST + k * 8, make_smem_desc_mn(da + k * 2048, FA_HALF_BYTES),
                                 idesc_mn, (it > 0 || k > 0) ? 1u : 0u);
                }
#pragma unroll
                for (int k = 0; k < FA_BM / UMMA_K; k++) {
                    // dK += dS^T Q
                    umma_bf16_ts(tmem_base + COL_DK, tmem_base + COL_DPT + k * 8, make_smem_desc_mn(qa + k * 2048, FA_HALF_BYTES),
                                 idesc_mn, (it > 0 || k > 0) ? 1u : 0u);
                }
                umma_commit(&bars->qdo_empty[stage]);
                if (++stage == 2) {
                    stage = 0;
                    phase ^= 1;
                }
                if (it + 1 < n_it) {
                    mbar_wait(&bars->qdo_full[stage], phase);
                    tc_fence_after();
                    issue_st_dpt(stage);
                } else {
                    umma_commit(&bars->d_full);
                }
            }
        }
    } else {
        const int tid = threadIdx.x - 64;  // 0..127
        const int lane_base = (warp & 3) * 32;
        const int key = j * FA_BN + lane_base + lane;  // key index of this thread's row
        const uint32_t lane_addr = (uint32_t)lane_base << 16;
        const uint32_t st_addr = tmem_base + lane_addr + COL_ST, dpt_addr = tmem_base + lane_addr + COL_DPT;
        for (int it = 0; it < n_it; it++) {
            const int h = hk * G + it / per_head, i = i0 + it % per_head;
            // per-query-row statistics of this (head, query block): thread `tid` fetches row `tid`
            float* st = &bars->stats[it & 1][0][0];
            {
                const int64_t sidx = ((int64_t)b * prm.H + h) * prm.S + i * FA_BM + tid;
                st[tid] = prm.lse[sidx] * 1.4426950408889634f;
                st[128 + tid] = prm.delta[sidx];
            }
            named_bar_sync(1, 128);
            mbar_wait(&bars->s_full, (uint32_t)(it & 1));
            tc_fence_after();
            const bool diag = prm.causal && i == j;
            const int q0 = i * FA_BM;
#pragma unroll 1
            for (int c = 0; c < 4; c++) {
                uint32_t s[32], dp[32], pp[16], ds[16];
                tmem_ld_32x32(st_addr + c * 32, s);
                tmem_ld_32x32(dpt_addr + c * 32, dp);
                tmem_wait_ld();
#pragma unroll
                for (int e = 0; e < 16; e++) {
                    const int qa = c * 32 + 2 * e;  // query row inside the block (columns of the transposed tiles)
                    const float2 l2 = *reinterpret_cast<const float2*>(&st[qa]);
                    const float2 dl = *reinterpret_cast<const float2*>(&st[128 + qa]);
                    float p0 = fast_exp2(fmaf(__uint_as_float(s[2 * e]), prm.scale_log2, -l2.x));
                    float p1 = fast_exp2(fmaf(__uint_as_float(s[2 * e + 1]), prm.scale_log2, -l2.y));
                    if (diag) {
                        if (key > q0 + qa) p0 = 0.f;
                        if (key > q0 + qa + 1) p1 = 0.f;
                    }
                    const float d0 = p0 * (__uint_as_float(dp[2 * e]) - dl.x) * prm.scale;
                    const float d1 = p1 * (__uint_as_float(dp[2 * e + 1]) - dl.y) * prm.scale;
                    __nv_bfloat162 hp = __floats2bfloat162_rn(p0, p1), hd = __floats2bfloat162_rn(d0, d1);
                    pp[e] = *reinterpret_cast<uint32_t*>(&hp);
                    ds[e] = *reinterpret_cast<uint32_t*>(&hd);
                }
                tmem_st_32x16(st_addr + c * 16, pp);   // P^T over the consumed S^T columns
                tmem_st_32x16(dpt_addr + c * 16, ds);  // dS^T over the consumed dP^T columns
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->p_full);
        }
        // epilogue: dV, dK -> bf16 rows of this key block
        mbar_wait(&bars->d_full, 0);
        tc_fence_after();
        __nv_bfloat16* vrow = prm.dv + (int64_t)b * prm.dv_batch_stride + (int64_t)key * prm.dv_row_stride + hk * FA_D;
        __nv_bfloat16* krow = prm.dk + (int64_t)b * prm.dk_batch_stride + (int64_t)key * prm.dk_row_stride + hk * FA_D;
#pragma unroll 1
        for (int t = 0; t < 2; t++) {
            __nv_bfloat16* orow = t == 0 ? vrow : krow;
            const uint32_t src = tmem_base + lane_addr + (t == 0 ? COL_DV : COL_DK);
#pragma unroll 1
            for (int c = 0; c < 4; c++) {
                uint32_t r[32];
                tmem_ld_32x32(src + c * 32, r);
                tmem_wait_ld();
#pragma unroll
                for (int v4 = 0; v4 < 4; v4++) {
                    uint4 pk;
                    __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                    for (int e = 0; e < 4; e++)
                        hh[e] = __floats2bfloat162_rn(__uint_as_float(r[v4 * 8 + 2 * e]), __uint_as_float(r[v4 * 8 + 2 * e + 1]));
                    *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = pk;
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

}  // namespace

extern "C" {

// All tensors bf16 [B, S, heads, 128] with element strides (last dim contiguous, heads packed 128 apart); lse [B, H, S]
// fp32 from the forward; dlse: gradient w.r.t. that log-sum-exp output ([B, H, S] fp32) or null; delta: fp32 scratch
// [B, H, S]. S must be a multiple of 128 (Sq == Skv; causal or full).
int dsb200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                          const float* dlse, float* delta, void* dq, void* dk, void* dv, int B, int S, int H, int Hkv,
                          int64_t q_row,
                          int64_t q_batch, int64_t k_row, int64_t k_batch, int64_t v_row, int64_t v_batch, int64_t o_row,
                          int64_t o_batch, int64_t do_row, int64_t do_batch, int64_t dq_row, int64_t dq_batch,
                          int64_t dk_row, int64_t dk_batch, int64_t dv_row, int64_t dv_batch, float scale, int causal,
                          cudaStream_t stream)
{
    if (B <= 0 || S <= 0) return 0;
    if (S % 128 != 0 || H % Hkv != 0) {
        dsb200_set_error("flash_attn_bwd: sequence length must be a multiple of 128 and H a multiple of Hkv");
        return -1;
    }
    const int64_t rows = (int64_t)B * S * H;
    fa_bwd_delta_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, stream>>>((const __nv_bfloat16*)o, (const __nv_bfloat16*)dout,
                                                                        dlse, delta, B, S, H, o_row, o_batch, do_row,
                                                                        do_batch);
    DSB_CHECK_LAUNCH("fa_bwd_delta");
    CUtensorMap mq, mdo, mk, mv;
    if (make_map_3d(&mq, q, B, S, (int64_t)H * FA_D, q_row, q_batch, 128) != 0) return -2;
    if (make_map_3d(&mdo, dout, B, S, (int64_t)H * FA_D, do_row, do_batch, 128) != 0) return -2;
    if (make_map_3d(&mk, k, B, S, (int64_t)Hkv * FA_D, k_row, k_batch, 128) != 0) return -2;
    if (make_map_3d(&mv, v, B, S, (int64_t)Hkv * FA_D, v_row, v_batch, 128) != 0) return -2;
    FbParams prm;
    prm.lse = lse, prm.delta = delta;
    prm.dq = (__nv_bfloat16*)dq, prm.dk = (__nv_bfloat16*)dk, prm.dv = (__nv_bfloat16*)dv;
    prm.B = B, prm.S = S, prm.H = H, prm.Hkv = Hkv;
    prm.dq_row_stride = dq_row, prm.dq_batch_stride = dq_batch;
    prm.dk_row_stride = dk_row, prm.dk_batch_stride = dk_batch;
    prm.dv_row_stride = dv_row, prm.dv_batch_stride = dv_batch;
    prm.scale = scale, prm.scale_log2 = scale * 1.4426950408889634f;
    prm.causal = causal, prm.nblk = S / 128;
    static bool attr_set[64] = {};
    int dev_ = 0;
    cudaGetDevice(&dev_);
    if (!attr_set[dev_ & 63]) {
        DSB_CHECK(cudaFuncSetAttribute(fa_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DQ_SMEM_BYTES));
        DSB_CHECK(cudaFuncSetAttribute(fa_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DKV_SMEM_BYTES));
        attr_set[dev_ & 63] = true;
    }
    fa_bwd_dkv_kernel<<<(unsigned)((int64_t)prm.nblk * B * Hkv), FB_THREADS, DKV_SMEM_BYTES, stream>>>(mq, mdo, mk, mv, prm);
    DSB_CHECK_LAUNCH("fa_bwd_dkv");
    fa_bwd_dq_kernel<<<(unsigned)((int64_t)prm.nblk * B * H), FB_THREADS, DQ_SMEM_BYTES, stream>>>(mq, mdo, mk, mv, prm);
    DSB_CHECK_LAUNCH("fa_bwd_dq");
    return 0;
}

}  // extern "C"
