// Flash-attention forward on the 5th-generation tensor cores (bf16, head_dim 128, optional causal mask, GQA).
//
//   O[b, i, h, :] = softmax_j(scale * Q[b, i, h, :] . K[b, j, h / (H/Hkv), :]) V[b, j, h / (H/Hkv), :]
//   LSE[b, h, i]  = log sum_j exp(scale * q.k)
//
// Operands stay in the projection layout [batch, seq, heads, 128] with arbitrary row / batch strides (so the views of
// a fused QKV projection are consumed in place); no transposes, no materialised scores.
//
// Reference: the reference has no attention kernel of its own for training; its models call the framework's SDPA /
// flash-attn (runtime/sequence_parallel/ulysses_sp.py:50, sequence/fpdt_layer.py:238-269), and its own fused softmax
// kernels (csrc/transformer/softmax_kernels.cu) materialise the [S, S] scores. This is the B200-native replacement.
//
// One CTA = one (batch, head) and TWO 128-row query tiles that ping-pong on the tensor pipe (while the softmax warps of
// one tile run, the MMAs of the other tile execute). 320 threads:
//   warp 0      TMA producer: Q tiles once, then K_j / V_j blocks (128 keys) through two 2-stage rings (3-D tensor
//               maps [batch, seq, heads*128], 128-byte swizzle, OOB rows zero-filled)
//   warp 1      MMA issuer (one elected lane): S_t = Q_t K_j^T (SS form, both operands K-major) into TMEM,
//               O_t += P_t V_j with P read straight from TMEM (TS form) and V as an MN-major smem operand
//   warps 2-5   softmax of tile 0, warps 6-9 softmax of tile 1: one thread per query row (TMEM lane == row, so row
//               max / row sum need no shuffles): tcgen05.ld S -> online softmax in the log2 domain -> bf16 P written
//               back over S with tcgen05.st -> lazy rescale of O in TMEM (only when the row max grew by > 2^8)
// TMEM (512 columns): S0 | S1 | O0 | O1, 128 fp32 columns each; P_t aliases the first 64 columns of S_t.
#include "attention_common.cuh"

namespace {

struct FaBars {
    uint64_t q_full;
    uint64_t k_full[FA_KV_STAGES], k_empty[FA_KV_STAGES];
    uint64_t v_full[FA_KV_STAGES], v_empty[FA_KV_STAGES];
    uint64_t s_full[2], p_full[2], o_full[2];
    uint32_t tmem_base;
};

// 2^x on the FMA/ALU pipes (Cody-Waite split + cubic minimax polynomial, relative error ~1e-4, far below the bf16
// rounding of P): the MUFU unit delivers 16 ex2/clk/SM, exactly the rate at which the tensor pipe consumes scores for
// head_dim 128, so a share of the exponentials is moved off it to give the softmax warps slack.
__device__ __forceinline__ float exp2_emu(float x)
{
    x = fmaxf(x, -126.f);
    const float t = __fadd_rd(x, 12582912.f);  // 1.5 * 2^23: floor(x) lands in the low mantissa bits
    const float f = x - (t - 12582912.f);      // fractional part in [0, 1)
    const float p = fmaf(fmaf(fmaf(0.077119f, f, 0.227564f), f, 0.695146f), f, 1.0f);
    return __uint_as_float(__float_as_uint(p) + (__float_as_uint(t) << 23));
}
// EMU_MASK: which of every 8 consecutive scores use exp2_emu (e.g. 0xA8 = positions 3, 5, 7 = 37.5 %)
template <unsigned EMU_MASK, int I>
__device__ __forceinline__ float exp2_sel(float x)
{
    if constexpr ((EMU_MASK >> (I & 7)) & 1u) return exp2_emu(x);
    else return fast_exp2(x);
}

struct FaParams {
    __nv_bfloat16* o;
    float* lse;  // [B, H, Sq]
    int B, Sq, Skv, H, Hkv;
    int64_t o_row_stride, o_batch_stride;
    float scale_log2;  // softmax scale * log2(e)
    int causal;
    int pairs;  // ceil(Sq / 256)
    // ragged / paged mode (FastGen prefill): `B` sequences packed back to back in q / o ([T_total, H, 128]); keys and
    // values come from a blocked cache [num_blocks, 2, 128, Hkv, 128] through a block table. Null in dense mode.
    const int* cu_q;         // [B] first packed row of each sequence
    const int* q_len;        // [B] query rows of each sequence in this batch
    const int* kv_len;       // [B] keys visible to the last query row (tokens already cached + q_len)
    const int* block_table;  // [B, max_blocks]
    int max_blocks;
    int lse_stride;          // elements between heads in lse (Sq dense, T_total ragged)
    // decode mode (one query token per "sequence", FastGen decode): the G = Hq / Hkv query heads that share a KV head
    // are the ROWS of the query tile (a [G, 128] box, zero-filled to 128 rows by TMA), the kernel's head index is the
    // KV head, and the key range is split over `pairs` CTAs whose normalised partial outputs + LSE are merged afterwards.
    int decode;              // 0 / 1
    int G;                   // query heads per KV head
    float* part_o;           // [T * Hq, splits, 128] fp32 (splits > 1)
    float* part_lse;         // [T * Hq, splits]
    // additive score biases (BIAS kernel variant; Evoformer attention): score = scale * q.k + bias1[b, col] +
    // bias2[b / bias2_div, h, row, col], both bf16, either may be null. Strides in elements.
    const __nv_bfloat16* bias1;
    const __nv_bfloat16* bias2;
    int64_t bias1_bstride, bias2_bstride, bias2_hstride, bias2_rstride;
    int bias2_div;
    float inv_scale;         // 1 / softmax scale: the bias is folded into the raw score as bias / scale
};

template <unsigned EMU_MASK, bool BIAS = false>
__global__ void __launch_bounds__(FA_THREADS, 1)
flash_attn_fwd_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                      const __grid_constant__ CUtensorMap map_v, const FaParams prm)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* q_smem = tiles;                                   // 2 tiles
    uint8_t* k_smem = tiles + 2 * FA_TILE_BYTES;               // FA_KV_STAGES
    uint8_t* v_smem = k_smem + FA_KV_STAGES * FA_TILE_BYTES;   // FA_KV_STAGES
    FaBars* bars = reinterpret_cast<FaBars*>(v_smem + FA_KV_STAGES * FA_TILE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // work item: heavy (late) query tiles first; consecutive CTAs share a KV head (L2 reuse)
    const int bh = prm.B * prm.H;
    const int pair = prm.pairs - 1 - (int)(blockIdx.x / bh);
    const int rem = (int)(blockIdx.x % bh);
    const int b = rem / prm.H, h = rem % prm.H;
    const int hk = h / (prm.H / prm.Hkv);
    const bool decode = prm.decode != 0;
    const int q0 = decode ? 0 : pair * 2 * FA_BM;  // first query row of tile 0 (relative to the sequence)
    const bool ragged = prm.cu_q != nullptr || decode;
    const int Sq = decode ? prm.G : (ragged ? prm.q_len[b] : prm.Sq);
    const int Skv = ragged ? prm.kv_len[b] : prm.Skv;
    const int row_base = decode ? 0 : (ragged ? prm.cu_q[b] : 0);  // packed row of the sequence's first query
    if (q0 >= Sq) return;                           // ragged batches are launched for the longest sequence
    // number of key blocks per tile
    const int nb_all = (Skv + FA_BN - 1) / FA_BN;
    // decode: this CTA's share of the key blocks (split index = pair)
    const int split_per = decode ? (nb_all + prm.pairs - 1) / prm.pairs : 0;
    const int jbase = decode ? min(nb_all, pair * split_per) : 0;
    int nb_t[2];
#pragma unroll
    for (int t = 0; t < 2; t++) {
        const int row0 = q0 + t * FA_BM;
        if (decode) {
            nb_t[t] = t == 0 ? min(nb_all, jbase + split_per) - jbase : 0;
        } else if (row0 >= Sq) {
            nb_t[t] = 0;
        } else if (prm.causal) {
            const int last_row = min(row0 + FA_BM, Sq) - 1 + (Skv - Sq);  // bottom-right aligned
            nb_t[t] = max(0, min(nb_all, last_row / FA_BN + 1));
        } else {
            nb_t[t] = nb_all;
        }
    }
    const int nb = max(nb_t[0], nb_t[1]);

    if (threadIdx.x == 0) {
        mbar_init(&bars->q_full, 1);
        for (int s = 0; s < FA_KV_STAGES; s++) {
            mbar_init(&bars->k_full[s], 1);
            mbar_init(&bars->k_empty[s], 1);
            mbar_init(&bars->v_full[s], 1);
            mbar_init(&bars->v_empty[s], 1);
        }
        for (int t = 0; t < 2; t++) {
            mbar_init(&bars->s_full[t], 1);
            mbar_init(&bars->p_full[t], 4);  // one arrive per softmax warp
            mbar_init(&bars->o_full[t], 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_k) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_v) : "memory");
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
    // column offsets
    constexpr uint32_t COL_S0 = 0, COL_O0 = 256;

    if (warp == 0) {
        // ============================ TMA producer ============================
        if (lane == 0 && nb > 0) {
            mbar_expect_tx(&bars->q_full, 2 * FA_TILE_BYTES);
#pragma unroll
            for (int t = 0; t < 2; t++)
#pragma unroll
                for (int hf = 0; hf < 2; hf++)
                    if (decode)  // q viewed as [T * Hkv, G, 128]: rows >= G of the box are zero-filled
                        tma_load_3d(q_smem + t * FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_q, &bars->q_full, hf * 64,
                                    t * FA_BM, b * prm.H + h);
                    else
                        tma_load_3d(q_smem + t * FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_q, &bars->q_full,
                                    h * FA_D + hf * 64, row_base + q0 + t * FA_BM, ragged ? 0 : b);
            int stage = 0;
            uint32_t phase = 0;
            for (int j = 0; j < nb; j++) {
                // dense: rows [j*128, +128) of batch b; paged: plane (block*2 + {K,V}) of the cache, rows [0, 128)
                const int blk = ragged ? prm.block_table[(int64_t)b * prm.max_blocks + jbase + j] : 0;
                const int krow = ragged ? 0 : j * FA_BN, kz = ragged ? blk * 2 : b, vz = ragged ? blk * 2 + 1 : b;
                mbar_wait(&bars->k_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->k_full[stage], FA_TILE_BYTES);
#pragma unroll
                for (int hf = 0; hf < 2; hf++)
                    tma_load_3d(k_smem + stage * FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_k, &bars->k_full[stage],
                                hk * FA_D + hf * 64, krow, kz);
                mbar_wait(&bars->v_empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->v_full[stage], FA_TILE_BYTES);
#pragma unroll
                for (int hf = 0; hf < 2; hf++)
                    tma_load_3d(v_smem + stage * FA_TILE_BYTES + hf * FA_HALF_BYTES, &map_v, &bars->v_full[stage],
                                hk * FA_D + hf * 64, krow, vz);
                if (++stage == FA_KV_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp == 1) {
        // ============================ MMA issuer ============================
        if (lane == 0 && nb > 0) {
            constexpr uint32_t idesc_qk = make_idesc_ex(FA_BM, FA_BN, 0, 0);  // S = Q K^T, both K-major
            constexpr uint32_t idesc_pv = make_idesc_ex(FA_BM, FA_D, 0, 1);   // O += P V, V is MN-major
            auto issue_qk = [&](int t, int kstage) {
                const uint32_t qa = smem_u32(q_smem + t * FA_TILE_BYTES), ka = smem_u32(k_smem + kstage * FA_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < FA_D / UMMA_K; k++) {
                    const uint32_t off = (uint32_t)(k >> 2) * FA_HALF_BYTES + (uint32_t)(k & 3) * 32;
                    umma_bf16(tmem_base + COL_S0 + t * 128, make_smem_desc(qa + off), make_smem_desc(ka + off), idesc_qk,
                              k > 0 ? 1u : 0u);
                }
                umma_commit(&bars->s_full[t]);
            };
            auto issue_pv = [&](int t, int vstage, bool first) {
                const uint32_t va = smem_u32(v_smem + vstage * FA_TILE_BYTES);
#pragma unroll
                for (int k = 0; k < FA_BN / UMMA_K; k++) {
                    // A: 16 bf16 of P per row = 8 TMEM columns per k step; B: 16 key rows = 2048 bytes per k step
                    umma_bf16_ts(tmem_base + COL_O0 + t * 128, tmem_base + COL_S0 + t * 128 + k * 8,
                                 make_smem_desc_mn(va + k * 2048, FA_HALF_BYTES), idesc_pv, (!first || k > 0) ? 1u : 0u);
                }
            };
            mbar_wait(&bars->q_full, 0);
            mbar_wait(&bars->k_full[0], 0);
            tc_fence_after();
            for (int t = 0; t < 2; t++)
                if (nb_t[t] > 0) issue_qk(t, 0);
            umma_commit(&bars->k_empty[0]);
            int vstage = 0;
            uint32_t vphase = 0;
            int kstage = 1 % FA_KV_STAGES;
            uint32_t kphase = (FA_KV_STAGES == 1) ? 1 : 0;
            for (int j = 0; j < nb; j++) {
                mbar_wait(&bars->v_full[vstage], vphase);
                const bool more = (j + 1 < nb);
                if (more) mbar_wait(&bars->k_full[kstage], kphase);
                tc_fence_after();
                for (int t = 0; t < 2; t++) {
                    if (j < nb_t[t]) {
                        mbar_wait(&bars->p_full[t], (uint32_t)(j & 1));
                        tc_fence_after();
                        issue_pv(t, vstage, j == 0);
                        if (j == nb_t[t] - 1) umma_commit(&bars->o_full[t]);
                        if (j + 1 < nb_t[t]) issue_qk(t, kstage);
                    }
                }
                umma_commit(&bars->v_empty[vstage]);
                if (++vstage == FA_KV_STAGES) {
                    vstage = 0;
                    vphase ^= 1;
                }
                if (more) {
                    umma_commit(&bars->k_empty[kstage]);
                    if (++kstage == FA_KV_STAGES) {
                        kstage = 0;
                        kphase ^= 1;
                    }
                }
            }
        }
    } else {
        // ============================ softmax / correction / epilogue ============================
        const int t = (warp - 2) >> 2;           // tile handled by this warp group
        const int lane_base = (warp & 3) * 32;   // TMEM lanes this warp may access
        const int row_in_tile = lane_base + lane;
        const int row = q0 + t * FA_BM + row_in_tile;  // query index
        const uint32_t lane_addr = (uint32_t)lane_base << 16;
        const uint32_t s_addr = tmem_base + lane_addr + COL_S0 + t * 128;
        const uint32_t o_addr = tmem_base + lane_addr + COL_O0 + t * 128;
        const int my_nb = nb_t[t];
        const int causal_off = Skv - Sq;
        float m_run = -INFINITY, l_run = 0.f;
        for (int j = 0; j < my_nb; j++) {
            mbar_wait(&bars->s_full[t], (uint32_t)(j & 1));
            tc_fence_after();
            const int col0 = (jbase + j) * FA_BN;
            // does this block need masking (diagonal block of the causal mask, or the tail of the key sequence)?
            const bool need_mask = (prm.causal && (col0 + FA_BN - 1 > q0 + t * FA_BM + causal_off)) || (col0 + FA_BN > Skv);
            const int col_limit = prm.causal ? min(Skv, row + causal_off + 1) : Skv;  // valid keys: col < col_limit
            // ---- the whole score row lives in registers: one TMEM round trip ----
            uint32_t r[4][32];
#pragma unroll
            for (int c = 0; c < 4; c++) tmem_ld_32x32(s_addr + c * 32, r[c]);
            tmem_wait_ld();
            float mx = -INFINITY;
            if constexpr (BIAS) {
                // s' = s + (bias1[col] + bias2[row, col]) / scale, so that s' * scale_log2 is the biased logit in the
                // log2 domain; 8 bf16 (16 bytes) per load, guarded at the sequence tail
                const bool row_ok = row < Sq;
                const __nv_bfloat16* b1 = prm.bias1 ? prm.bias1 + (int64_t)b * prm.bias1_bstride : nullptr;
                const __nv_bfloat16* b2 = (prm.bias2 && row_ok)
                    ? prm.bias2 + (int64_t)(b / prm.bias2_div) * prm.bias2_bstride + (int64_t)h * prm.bias2_hstride +
                          (int64_t)row * prm.bias2_rstride
                    : nullptr;
                const bool vec_ok = (col0 + FA_BN <= Skv) && (Skv % 8 == 0);
#pragma unroll
                for (int c = 0; c < 4; c++) {
#pragma unroll
                    for (int i8 = 0; i8 < 32; i8 += 8) {
                        const int col = col0 + c * 32 + i8;
                        float add[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                        if (vec_ok) {
                            if (b1) {
                                const uint4 raw = *reinterpret_cast<const uint4*>(b1 + col);
                                const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
                                for (int q4 = 0; q4 < 4; q4++) {
                                    const float2 f = __bfloat1622float2(hp[q4]);
                                    add[2 * q4] += f.x, add[2 * q4 + 1] += f.y;
                                }
                            }
                            if (b2) {
                                const uint4 raw = *reinterpret_cast<const uint4*>(b2 + col);
                                const __nv_bfloat162* hp = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
                                for (int q4 = 0; q4 < 4; q4++) {
                                    const float2 f = __bfloat1622float2(hp[q4]);
                                    add[2 * q4] += f.x, add[2 * q4 + 1] += f.y;
                                }
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 8; e++) {
                                if (col + e < Skv) {
                                    if (b1) add[e] += __bfloat162float(b1[col + e]);
                                    if (b2) add[e] += __bfloat162float(b2[col + e]);
                                }
                            }
                        }
#pragma unroll
                        for (int e = 0; e < 8; e++)
                            r[c][i8 + e] = __float_as_uint(fmaf(add[e], prm.inv_scale, __uint_as_float(r[c][i8 + e])));
                    }
                }
            }
            if (need_mask) {
#pragma unroll
                for (int c = 0; c < 4; c++)
#pragma unroll
                    for (int i = 0; i < 32; i++)
                        if (col0 + c * 32 + i >= col_limit) r[c][i] = 0xff800000u;  // -inf
            }
            {
                float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // 4 independent chains
#pragma unroll
                for (int c = 0; c < 4; c++)
#pragma unroll
                    for (int i = 0; i < 32; i++) mx4[i & 3] = fmaxf(mx4[i & 3], __uint_as_float(r[c][i]));
                mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
            }
            mx *= prm.scale_log2;  // scale > 0: max commutes with the scaling
            float m_new = fmaxf(m_run, mx);
            // lazy rescale: keep the stale max while the new one is within 2^8 of it (p stays <= 256)
            if (m_run != -INFINITY && m_new - m_run <= 8.f) m_new = m_run;
            const float alpha = (m_run == -INFINITY) ? 0.f : fast_exp2(m_run - m_new);
            const float m_use = (m_new == -INFINITY) ? 0.f : m_new;  // fully masked row so far: avoid inf - inf
            // ---- p = exp2(s * scale - m) (masked entries: exp2(-inf) = 0), bf16 P over the S columns ----
            float rs4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int c = 0; c < 4; c++) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                    // 8 scores per step: compile-time choice of MUFU vs FMA-pipe exponential per position
                    const float x0 = fmaf(__uint_as_float(r[c][2 * i + 0]), prm.scale_log2, -m_use);
                    const float x1 = fmaf(__uint_as_float(r[c][2 * i + 1]), prm.scale_log2, -m_use);
                    const float x2 = fmaf(__uint_as_float(r[c][2 * i + 2]), prm.scale_log2, -m_use);
                    const float x3 = fmaf(__uint_as_float(r[c][2 * i + 3]), prm.scale_log2, -m_use);
                    const float x4 = fmaf(__uint_as_float(r[c][2 * i + 4]), prm.scale_log2, -m_use);
                    const float x5 = fmaf(__uint_as_float(r[c][2 * i + 5]), prm.scale_log2, -m_use);
                    const float x6 = fmaf(__uint_as_float(r[c][2 * i + 6]), prm.scale_log2, -m_use);
                    const float x7 = fmaf(__uint_as_float(r[c][2 * i + 7]), prm.scale_log2, -m_use);
                    const float p0 = exp2_sel<EMU_MASK, 0>(x0), p1 = exp2_sel<EMU_MASK, 1>(x1);
                    const float p2 = exp2_sel<EMU_MASK, 2>(x2), p3 = exp2_sel<EMU_MASK, 3>(x3);
                    const float p4 = exp2_sel<EMU_MASK, 4>(x4), p5 = exp2_sel<EMU_MASK, 5>(x5);
                    const float p6 = exp2_sel<EMU_MASK, 6>(x6), p7 = exp2_sel<EMU_MASK, 7>(x7);
                    rs4[0] += p0 + p1;
                    rs4[1] += p2 + p3;
                    rs4[2] += p4 + p5;
                    rs4[3] += p6 + p7;
                    __nv_bfloat162 h0 = __floats2bfloat162_rn(p0, p1), h1 = __floats2bfloat162_rn(p2, p3);
                    __nv_bfloat162 h2 = __floats2bfloat162_rn(p4, p5), h3 = __floats2bfloat162_rn(p6, p7);
                    pk[i + 0] = *reinterpret_cast<uint32_t*>(&h0);
                    pk[i + 1] = *reinterpret_cast<uint32_t*>(&h1);
                    pk[i + 2] = *reinterpret_cast<uint32_t*>(&h2);
                    pk[i + 3] = *reinterpret_cast<uint32_t*>(&h3);
                }
                tmem_st_32x16(s_addr + c * 16, pk);
            }
            const float rowsum = (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
            l_run = l_run * alpha + rowsum;
            m_run = m_new;
            // ---- rescale O (the previous P V has completed: s_full of this block was committed after it) ----
            if (j > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll 1
                for (int c = 0; c < 4; c++) {
                    uint32_t r[32];
                    tmem_ld_32x32(o_addr + c * 32, r);
                    tmem_wait_ld();
#pragma unroll
                    for (int i = 0; i < 32; i++) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                    tmem_st_32x32(o_addr + c * 32, r);
                }
            }
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->p_full[t]);
        }
        // ---- epilogue: O / l -> bf16 -> global, LSE ----
        if (my_nb > 0) {
            mbar_wait(&bars->o_full[t], 0);
            tc_fence_after();
            const float inv_l = (l_run > 0.f) ? 1.f / l_run : 0.f;
            const bool valid = row < Sq;
            // decode: row = query head inside the KV-head group; o is [T, Hq, 128] and the kernel's H is Hkv
            const int64_t head_row = decode ? ((int64_t)b * prm.H + h) * prm.G + row : 0;
            __nv_bfloat16* orow = decode ? prm.o + head_row * FA_D
                                         : prm.o + (ragged ? 0 : (int64_t)b * prm.o_batch_stride) +
                                               (int64_t)(row_base + row) * prm.o_row_stride + h * FA_D;
            const bool partial = decode && prm.pairs > 1;
            float* prow = partial ? prm.part_o + (head_row * prm.pairs + pair) * FA_D : nullptr;
#pragma unroll 1
            for (int c = 0; c < 4; c++) {
                uint32_t r[32];
                tmem_ld_32x32(o_addr + c * 32, r);
                tmem_wait_ld();
                if (valid && partial) {
#pragma unroll
                    for (int v4 = 0; v4 < 8; v4++)
                        *reinterpret_cast<float4*>(prow + c * 32 + v4 * 4) =
                            make_float4(__uint_as_float(r[v4 * 4]) * inv_l, __uint_as_float(r[v4 * 4 + 1]) * inv_l,
                                        __uint_as_float(r[v4 * 4 + 2]) * inv_l, __uint_as_float(r[v4 * 4 + 3]) * inv_l);
                } else if (valid) {
#pragma unroll
                    for (int v4 = 0; v4 < 4; v4++) {
                        uint4 pk;
                        __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                        for (int i = 0; i < 4; i++)
                            hh[i] = __floats2bfloat162_rn(__uint_as_float(r[v4 * 8 + 2 * i]) * inv_l,
                                                          __uint_as_float(r[v4 * 8 + 2 * i + 1]) * inv_l);
                        *reinterpret_cast<uint4*>(orow + c * 32 + v4 * 8) = pk;
                    }
                }
            }
            if (valid && partial) {
                prm.part_lse[head_row * prm.pairs + pair] = (l_run > 0.f) ? (m_run + log2f(l_run)) : -INFINITY;  // log2
            } else if (valid && prm.lse != nullptr && !decode) {
                const float lse = (l_run > 0.f) ? (m_run + log2f(l_run)) * 0.6931471805599453f : -INFINITY;
                prm.lse[((int64_t)(ragged ? 0 : b) * prm.H + h) * prm.lse_stride + row_base + row] = lse;
            }
        } else if (decode) {
            // a split beyond the end of this sequence's keys: an empty partial (weight 0 in the merge)
            if (t == 0 && row < Sq && prm.pairs > 1)
                prm.part_lse[(((int64_t)b * prm.H + h) * prm.G + row) * prm.pairs + pair] = -INFINITY;
        } else if (row < Sq && q0 + t * FA_BM < Sq) {
            // no visible key at all (causal with Skv < Sq): zeros
            __nv_bfloat16* orow = prm.o + (ragged ? 0 : (int64_t)b * prm.o_batch_stride) +
                                  (int64_t)(row_base + row) * prm.o_row_stride + h * FA_D;
            for (int i = 0; i < FA_D; i += 8) *reinterpret_cast<uint4*>(orow + i) = make_uint4(0, 0, 0, 0);
            if (prm.lse != nullptr)
                prm.lse[((int64_t)(ragged ? 0 : b) * prm.H + h) * prm.lse_stride + row_base + row] = -INFINITY;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
    }
}

// decode split-KV merge: out[row] = sum_s w_s * part_o[row, s] with w_s = 2^(lse_s - max) / sum 2^(lse - max)
__global__ void __launch_bounds__(128) flash_decode_merge_kernel(const float* __restrict__ part_o,
                                                                 const float* __restrict__ part_lse,
                                                                 __nv_bfloat16* __restrict__ out, int splits)
{
    const int64_t row = blockIdx.x;
    const int d = threadIdx.x;
    float mx = -INFINITY;
    for (int s = 0; s < splits; s++) mx = fmaxf(mx, part_lse[row * splits + s]);
    float den = 0.f, acc = 0.f;
    for (int s = 0; s < splits; s++) {
        const float l = part_lse[row * splits + s];
        if (l == -INFINITY) continue;
        const float w = exp2f(l - mx);
        den += w;
        acc += w * part_o[(row * splits + s) * FA_D + d];
    }
    out[row * FA_D + d] = __float2bfloat16_rn(den > 0.f ? acc / den : 0.f);
}

}  // namespace

namespace {

int launch_flash_fwd(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const FaParams& prm,
                     cudaStream_t stream)
{
    // share of exponentials computed on the FMA pipes instead of the MUFU (tuning knob DSB200_FA_EMU = 0..3 ->
    // 0 / 25 / 37.5 / 50 %). Measured on B200 (S = 4096..8192): 0 % is fastest (1.24 PFLOP/s non-causal at 8192; every
    // extra 12.5 % costs ~3 %), the softmax warps are issue-bound before they are MUFU-bound, so the default is 0.
    static int emu = -1;
    if (emu < 0) {
        const char* e = getenv("DSB200_FA_EMU");
        emu = e ? atoi(e) : 0;
        if (emu < 0 || emu > 3) emu = 0;
    }
    const int64_t grid = (int64_t)prm.pairs * prm.B * prm.H;
    int dev_ = 0;
    cudaGetDevice(&dev_);
#define FA_LAUNCH(MASK)                                                                                             \
    do {                                                                                                            \
        static bool attr_set[64] = {};                                                                              \
        if (!attr_set[dev_ & 63]) {                                                                                 \
            DSB_CHECK(cudaFuncSetAttribute(flash_attn_fwd_kernel<MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                           FA_SMEM_BYTES));                                                         \
            attr_set[dev_ & 63] = true;                                                                             \
        }                                                                                                           \
        flash_attn_fwd_kernel<MASK><<<(unsigned)grid, FA_THREADS, FA_SMEM_BYTES, stream>>>(mq, mk, mv, prm);        \
    } while (0)
    if (emu == 0)
        FA_LAUNCH(0x00u);
    else if (emu == 1)
        FA_LAUNCH(0x88u);  // 25 %
    else if (emu == 2)
        FA_LAUNCH(0xA8u);  // 37.5 %
    else
        FA_LAUNCH(0xAAu);  // 50 %
#undef FA_LAUNCH
    DSB_CHECK_LAUNCH("flash_attn_fwd");
    return 0;
}

int launch_flash_fwd_bias(const CUtensorMap& mq, const CUtensorMap& mk, const CUtensorMap& mv, const FaParams& prm,
                          cudaStream_t stream)
{
    const int64_t grid = (int64_t)prm.pairs * prm.B * prm.H;
    int dev_ = 0;
    cudaGetDevice(&dev_);
    static bool attr_set[64] = {};
    if (!attr_set[dev_ & 63]) {
        DSB_CHECK(cudaFuncSetAttribute(flash_attn_fwd_kernel<0x00u, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       FA_SMEM_BYTES));
        attr_set[dev_ & 63] = true;
    }
    flash_attn_fwd_kernel<0x00u, true><<<(unsigned)grid, FA_THREADS, FA_SMEM_BYTES, stream>>>(mq, mk, mv, prm);
    DSB_CHECK_LAUNCH("flash_attn_fwd_bias");
    return 0;
}

}  // namespace

extern "C" {

// q [B, Sq, H, 128], k / v [B, Skv, Hkv, 128], o [B, Sq, H, 128] (all bf16; strides in elements, last dim contiguous,
// heads packed 128 apart), lse [B, H, Sq] fp32 (may be null). causal: bottom-right aligned when Skv != Sq.
int dsb200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int Sq, int Skv,
                          int H, int Hkv, int64_t q_row_stride, int64_t q_batch_stride, int64_t k_row_stride,
                          int64_t k_batch_stride, int64_t v_row_stride, int64_t v_batch_stride, int64_t o_row_stride,
                          int64_t o_batch_stride, float scale, int causal, cudaStream_t stream)
{
    if (B <= 0 || Sq <= 0 || Skv <= 0) return 0;
    if (H % Hkv != 0 || (o_row_stride % 8) || (o_batch_stride % 8) || ((uintptr_t)o & 15)) {
        dsb200_set_error("flash_attn_fwd: H must be a multiple of Hkv; output must be 16-byte aligned");
        return -1;
    }
    CUtensorMap mq, mk, mv;
    if (make_map_3d(&mq, q, B, Sq, (int64_t)H * FA_D, q_row_stride, q_batch_stride, FA_BM) != 0) return -2;
    if (make_map_3d(&mk, k, B, Skv, (int64_t)Hkv * FA_D, k_row_stride, k_batch_stride, FA_BN) != 0) return -2;
    if (make_map_3d(&mv, v, B, Skv, (int64_t)Hkv * FA_D, v_row_stride, v_batch_stride, FA_BN) != 0) return -2;
    FaParams prm = {};
    prm.o = (__nv_bfloat16*)o;
    prm.lse = lse;
    prm.B = B, prm.Sq = Sq, prm.Skv = Skv, prm.H = H, prm.Hkv = Hkv;
    prm.o_row_stride = o_row_stride, prm.o_batch_stride = o_batch_stride;
    prm.scale_log2 = scale * 1.4426950408889634f;
    prm.causal = causal;
    prm.pairs = (Sq + 2 * FA_BM - 1) / (2 * FA_BM);
    prm.lse_stride = Sq;
    return launch_flash_fwd(mq, mk, mv, prm, stream);
}

// Attention with additive score biases (Evoformer / any bias that broadcasts this way), non-causal:
//   score[b, h, i, j] = scale * q[b, i, h, :] . k[b, j, h, :] + bias1[b, j] + bias2[b / bias2_div, h, i, j]
// q / k / v / o as in dsb200_flash_attn_fwd (head_dim 128: pad smaller heads with zeros); bias1 [B, Skv] (row stride
// bias1_bstride), bias2 [B / bias2_div, H, Sq, Skv] (strides given), both bf16, either may be null. For 16-byte bias
// loads Skv should be a multiple of 8 and the bias rows 16-byte aligned (otherwise a scalar path is taken).
int dsb200_flash_attn_fwd_bias(const void* q, const void* k, const void* v, void* o, float* lse, const void* bias1,
                               const void* bias2, int B, int Sq, int Skv, int H, int64_t q_row_stride,
                               int64_t q_batch_stride, int64_t k_row_stride, int64_t k_batch_stride,
                               int64_t v_row_stride, int64_t v_batch_stride, int64_t o_row_stride,
                               int64_t o_batch_stride, int64_t bias1_bstride, int64_t bias2_bstride,
                               int64_t bias2_hstride, int64_t bias2_rstride, int bias2_div, float scale,
                               cudaStream_t stream)
{
    if (B <= 0 || Sq <= 0 || Skv <= 0) return 0;
    if ((o_row_stride % 8) || (o_batch_stride % 8) || ((uintptr_t)o & 15) || scale <= 0.f || bias2_div < 1) {
        dsb200_set_error("flash_attn_fwd_bias: output must be 16-byte aligned, scale > 0, bias2_div >= 1");
        return -1;
    }
    const bool aligned = !(((uintptr_t)bias1 | (uintptr_t)bias2) & 15) && (bias1_bstride % 8 == 0) &&
                         (bias2_bstride % 8 == 0) && (bias2_hstride % 8 == 0) && (bias2_rstride % 8 == 0);
    CUtensorMap mq, mk, mv;
    if (make_map_3d(&mq, q, B, Sq, (int64_t)H * FA_D, q_row_stride, q_batch_stride, FA_BM) != 0) return -2;
    if (make_map_3d(&mk, k, B, Skv, (int64_t)H * FA_D, k_row_stride, k_batch_stride, FA_BN) != 0) return -2;
    if (make_map_3d(&mv, v, B, Skv, (int64_t)H * FA_D, v_row_stride, v_batch_stride, FA_BN) != 0) return -2;
    FaParams prm = {};
    prm.o = (__nv_bfloat16*)o;
    prm.lse = lse;
    prm.B = B, prm.Sq = Sq, prm.H = H, prm.Hkv = H;
    // a Skv that is not a multiple of 8, or unaligned bias rows, must not take the 16-byte path: the kernel keys that
    // on Skv % 8, so report an odd Skv to it only through the scalar-path condition below
    prm.Skv = Skv;
    prm.o_row_stride = o_row_stride, prm.o_batch_stride = o_batch_stride;
    prm.scale_log2 = scale * 1.4426950408889634f;
    prm.inv_scale = 1.0f / scale;
    prm.causal = 0;
    prm.pairs = (Sq + 2 * FA_BM - 1) / (2 * FA_BM);
    prm.lse_stride = Sq;
    prm.bias1 = (const __nv_bfloat16*)bias1, prm.bias2 = (const __nv_bfloat16*)bias2;
    prm.bias1_bstride = bias1_bstride, prm.bias2_bstride = bias2_bstride;
    prm.bias2_hstride = bias2_hstride, prm.bias2_rstride = bias2_rstride, prm.bias2_div = bias2_div;
    if (!aligned) {
        dsb200_set_error("flash_attn_fwd_bias: bias pointers / strides must be 16-byte aligned (multiples of 8 elements)");
        return -1;
    }
    return launch_flash_fwd_bias(mq, mk, mv, prm, stream);
}

// Ragged prefill over a blocked KV cache (FastGen): `n_seqs` sequences packed back to back in q / o [T_total, H, 128]
// (row stride in elements); sequence s owns rows [cu_q[s], cu_q[s] + q_len[s]) and attends causally to the first
// kv_len[s] keys of its blocks (kv_len - q_len of them were cached by earlier chunks). `kv` is one layer of the cache,
// [num_blocks, 2, 128, Hkv, 128] -- the block size must equal the kernel's key block (128). cu_q / q_len / kv_len /
// block_table are device int32 arrays; max_q_len is the host-side maximum of q_len. lse [H, T_total] may be null.
// Reference: inference/v2/kernels/ragged_ops/blocked_flash (paged FA2 fork).
int dsb200_flash_attn_fwd_paged(const void* q, const void* kv, void* o, float* lse, const int* cu_q, const int* q_len,
                                const int* kv_len, const int* block_table, int n_seqs, int T_total, int max_q_len,
                                int H, int Hkv, int num_blocks, int max_blocks, int64_t q_row_stride,
                                int64_t o_row_stride, float scale, cudaStream_t stream)
{
    if (n_seqs <= 0 || T_total <= 0 || max_q_len <= 0) return 0;
    if (H % Hkv != 0 || (o_row_stride % 8) || ((uintptr_t)o & 15)) {
        dsb200_set_error("flash_attn_fwd_paged: H must be a multiple of Hkv; output must be 16-byte aligned");
        return -1;
    }
    CUtensorMap mq, mk;
    if (make_map_3d(&mq, q, 1, T_total, (int64_t)H * FA_D, q_row_stride, (int64_t)T_total * q_row_stride, FA_BM) != 0)
        return -2;
    // the cache as [num_blocks * 2 planes, 128 rows, Hkv * 128]
    const int64_t plane = (int64_t)FA_BN * Hkv * FA_D;
    if (make_map_3d(&mk, kv, (int64_t)num_blocks * 2, FA_BN, (int64_t)Hkv * FA_D, (int64_t)Hkv * FA_D, plane, FA_BN) != 0)
        return -2;
    FaParams prm = {};
    prm.o = (__nv_bfloat16*)o;
    prm.lse = lse;
    prm.B = n_seqs, prm.Sq = max_q_len, prm.Skv = 0, prm.H = H, prm.Hkv = Hkv;
    prm.o_row_stride = o_row_stride, prm.o_batch_stride = 0;
    prm.scale_log2 = scale * 1.4426950408889634f;
    prm.causal = 1;
    prm.pairs = (max_q_len + 2 * FA_BM - 1) / (2 * FA_BM);
    prm.cu_q = cu_q, prm.q_len = q_len, prm.kv_len = kv_len, prm.block_table = block_table;
    prm.max_blocks = max_blocks;
    prm.lse_stride = T_total;
    return launch_flash_fwd(mq, mk, mk, prm, stream);
}


// Decode over a blocked KV cache on the tcgen05 kernel: T tokens (one per sequence), q [T, Hq, 128] contiguous,
// kv one cache layer [num_blocks, 2, 128, Hkv, 128], kv_len[T] keys visible to each token, block_table [T, max_blocks].
// splits > 1 needs part_o (T*Hq*splits*128 floats) and part_lse (T*Hq*splits floats). o [T, Hq, 128] bf16.
// Reference: inference/v2/kernels/ragged_ops/blocked_flash (decode path).
int dsb200_flash_decode_paged(const void* q, const void* kv, void* o, const int* kv_len, const int* block_table,
                              float* part_o, float* part_lse, int T, int Hq, int Hkv, int num_blocks, int max_blocks,
                              int splits, float scale, cudaStream_t stream)
{
    if (T <= 0) return 0;
    const int G = Hkv > 0 ? Hq / Hkv : 0;
    if (Hkv <= 0 || Hq % Hkv != 0 || G > FA_BM) {
        dsb200_set_error("flash_decode_paged: Hq must be a multiple of Hkv (at most 128 query heads per KV head)");
        return -1;
    }
    if (splits < 1) splits = 1;
    if (splits > 1 && (part_o == nullptr || part_lse == nullptr)) {
        dsb200_set_error("flash_decode_paged: split-KV needs the partial-result buffers");
        return -1;
    }
    CUtensorMap mq, mk;
    // q as [T * Hkv, G, 128]: one [G, 128] box per (token, KV head); rows G..127 of the 128-row box are OOB -> zeros
    if (make_map_3d(&mq, q, (int64_t)T * Hkv, G, FA_D, FA_D, (int64_t)G * FA_D, FA_BM) != 0) return -2;
    const int64_t plane = (int64_t)FA_BN * Hkv * FA_D;
    if (make_map_3d(&mk, kv, (int64_t)num_blocks * 2, FA_BN, (int64_t)Hkv * FA_D, (int64_t)Hkv * FA_D, plane, FA_BN) != 0)
        return -2;
    FaParams prm = {};
    prm.o = (__nv_bfloat16*)o;
    prm.B = T, prm.Sq = G, prm.Skv = 0, prm.H = Hkv, prm.Hkv = Hkv;
    prm.o_row_stride = FA_D, prm.o_batch_stride = 0;
    prm.scale_log2 = scale * 1.4426950408889634f;
    prm.causal = 0;
    prm.pairs = splits;
    prm.kv_len = kv_len, prm.block_table = block_table, prm.max_blocks = max_blocks;
    prm.decode = 1, prm.G = G;
    prm.part_o = part_o, prm.part_lse = part_lse;
    int rc = launch_flash_fwd(mq, mk, mk, prm, stream);
    if (rc != 0) return rc;
    if (splits > 1) {
        flash_decode_merge_kernel<<<T * Hq, 128, 0, stream>>>(part_o, part_lse, (__nv_bfloat16*)o, splits);
        DSB_CHECK_LAUNCH("flash_decode_merge");
    }
    return 0;
}

}  // extern "C"
