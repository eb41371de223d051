// FastGen ragged-batch kernels over a blocked (paged) KV cache.
//
// Cache layout of one layer: [num_blocks, 2 (K | V), block_size, kv_heads, head_dim] -- the K (or V) plane of a block
// is one contiguous [block_size, kv_heads * head_dim] matrix, so a decode CTA streams it with 16-byte loads and the
// tcgen05 prefill kernel can address it with one TMA box per block.
//
//   dsb200_ragged_kv_append_rope   RoPE (half rotation, optional partial rotary) on q in place and on k, then k / v of
//                                  every token of the ragged batch are written to their cache slots
//   dsb200_ragged_paged_attention  every token attends to keys [0, pos] of its own sequence through the block table;
//                                  one CTA per (token, kv head [, split of the key range]); the G = Hq / Hkv query
//                                  heads that share a KV head are processed together so K / V are read once;
//                                  online softmax in the log2 domain; split-KV partials merged by a second kernel
//
// Reference: inference/v2/kernels/ragged_ops/linear_blocked_kv_rotary/blocked_kv_rotary_cuda.cu (rotary + KV copy),
// ragged_ops/blocked_flash (paged flash attention, an external FA2 fork the reference links as -lblockedflash),
// ragged_ops/atom_builder (work decomposition). The reference kernels are Ampere SIMT / mma.sync; decode attention is
// HBM-bound, so this one is SIMT as well and sized to stream K/V at memory speed on 148 SMs.
#include "common.cuh"

namespace {

constexpr int PA_THREADS = 128;
constexpr int PA_WARPS = PA_THREADS / 32;
constexpr int PA_TILE = 32;  // key positions per warp tile

struct PagedArgs {
    const __nv_bfloat16* q;   // [T, Hq, D], token stride q_stride
    const __nv_bfloat16* kv;  // layer base of the cache
    __nv_bfloat16* out;       // [T, Hq, D] contiguous
    const int* tok_seq;       // [T] sequence slot of each token (row of block_table)
    const int* tok_pos;       // [T] absolute position of each token inside its sequence
    const int* block_table;   // [n_seqs, max_blocks]
    const int* tok_row;       // [T] row of q / out for each listed token (null: row == index)
    float* part_acc;          // [T, Hq, splits, D] (splits > 1)
    float* part_ml;           // [T, Hq, splits, 2]
    int64_t q_stride;
    int T, Hq, Hkv, block_size, max_blocks, splits;
    float scale_log2;
};

template <int D>
__device__ __forceinline__ void load_row_to_smem(__nv_bfloat16* dst, const __nv_bfloat16* src, int lane_in_row)
{
    // one 16-byte piece of a D-element row
    *reinterpret_cast<uint4*>(dst + lane_in_row * 8) = __ldg(reinterpret_cast<const uint4*>(src + lane_in_row * 8));
}

// D: head dim (64 or 128). G: query heads per kv head handled by this CTA (1, 2, 4, 8).
template <int D, int G>
__global__ void __launch_bounds__(PA_THREADS) paged_attention_kernel(const PagedArgs a)
{
    constexpr int ROW = D + 8;                // padded smem row (bf16 elements): 16-byte phases stay conflict-free
    constexpr int LANES_PER_ROW = D / 8;      // 16-byte pieces per row
    constexpr int ROWS_PER_LOAD = 32 / LANES_PER_ROW;
    constexpr int DV = D / 32;                // output columns owned by a lane
    constexpr int N_LOADS = PA_TILE / ROWS_PER_LOAD;
    // one tile buffer per warp: holds the K tile for the scores, then the V tile (prefetched into registers while the
    // scores are computed) for P V; after the loop it carries the warp's partial output
    __shared__ __align__(16) __nv_bfloat16 s_kv[PA_WARPS][PA_TILE * ROW];
    __shared__ __align__(16) float s_q[G * D];
    __shared__ __align__(16) float s_p[PA_WARPS][G * PA_TILE];
    __shared__ float s_m[PA_WARPS][G], s_l[PA_WARPS][G];
    static_assert(G * D * 4 <= PA_TILE * ROW * 2, "partial output must fit in the tile buffer");

    const int ti = blockIdx.x, hk = blockIdx.y, split = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int seq = a.tok_seq[ti], pos = a.tok_pos[ti];
    const int t = a.tok_row != nullptr ? a.tok_row[ti] : ti;  // row of q / out
    const int n_keys = pos + 1;
    const int n_tiles = (n_keys + PA_TILE - 1) / PA_TILE;
    // this split's tile range
    const int per = (n_tiles + a.splits - 1) / a.splits;
    const int tile_lo = split * per, tile_hi = min(n_tiles, tile_lo + per);

    // stage q (scaled into the log2 domain) as fp32
    for (int i = threadIdx.x; i < G * D; i += PA_THREADS) {
        const int g = i / D, d = i % D;
        s_q[i] = __bfloat162float(a.q[(int64_t)t * a.q_stride + (int64_t)(hk * G + g) * D + d]) * a.scale_log2;
    }
    __syncthreads();

    float m[G], l[G], acc[G][DV];
#pragma unroll
    for (int g = 0; g < G; g++) {
        m[g] = -INFINITY;
        l[g] = 0.f;
#pragma unroll
        for (int c = 0; c < DV; c++) acc[g][c] = 0.f;
    }
    const int64_t plane = (int64_t)a.block_size * a.Hkv * D;  // elements of one K or V plane
    const int* table = a.block_table + (int64_t)seq * a.max_blocks;
    __nv_bfloat16* wk = s_kv[warp];
    __nv_bfloat16* wv = s_kv[warp];
    float* wp = s_p[warp];

    for (int tile = tile_lo + warp; tile < tile_hi; tile += PA_WARPS) {
        const int key0 = tile * PA_TILE;
        // ---- global -> shared: 32 K rows; the 32 V rows wait in registers (rows beyond n_keys are zero-filled) ----
        uint4 vreg[N_LOADS];
#pragma unroll
        for (int i = 0; i < N_LOADS; i++) {
            const int r = i * ROWS_PER_LOAD + lane / LANES_PER_ROW, piece = lane % LANES_PER_ROW;
            const int key = key0 + r;
            uint4 kk = make_uint4(0, 0, 0, 0);
            vreg[i] = make_uint4(0, 0, 0, 0);
            if (key < n_keys) {
                const int blk = table[key / a.block_size], off = key % a.block_size;
                const __nv_bfloat16* kp = a.kv + ((int64_t)blk * 2) * plane + ((int64_t)off * a.Hkv + hk) * D + piece * 8;
                kk = __ldg(reinterpret_cast<const uint4*>(kp));
                vreg[i] = __ldg(reinterpret_cast<const uint4*>(kp + plane));
            }
            *reinterpret_cast<uint4*>(wk + r * ROW + piece * 8) = kk;
        }
        __syncwarp();
        // ---- scores: lane = key position ----
        float s[G];
#pragma unroll
        for (int g = 0; g < G; g++) s[g] = 0.f;
#pragma unroll 4
        for (int c = 0; c < D / 8; c++) {
            const uint4 raw = *reinterpret_cast<const uint4*>(wk + lane * ROW + c * 8);
            const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
            float kf[8];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const float2 f = __bfloat1622float2(h2[i]);
                kf[2 * i] = f.x;
                kf[2 * i + 1] = f.y;
            }
#pragma unroll
            for (int g = 0; g < G; g++) {
                const float4 q0 = *reinterpret_cast<const float4*>(s_q + g * D + c * 8);
                const float4 q1 = *reinterpret_cast<const float4*>(s_q + g * D + c * 8 + 4);
                s[g] = fmaf(kf[0], q0.x, s[g]);
                s[g] = fmaf(kf[1], q0.y, s[g]);
                s[g] = fmaf(kf[2], q0.z, s[g]);
                s[g] = fmaf(kf[3], q0.w, s[g]);
                s[g] = fmaf(kf[4], q1.x, s[g]);
                s[g] = fmaf(kf[5], q1.y, s[g]);
                s[g] = fmaf(kf[6], q1.z, s[g]);
                s[g] = fmaf(kf[7], q1.w, s[g]);
            }
        }
        const bool valid = key0 + lane < n_keys;
        // ---- online softmax (per query head), probabilities to shared memory ----
        float alpha[G];
#pragma unroll
        for (int g = 0; g < G; g++) {
            const float sv = valid ? s[g] : -INFINITY;
            const float mt = dsb::warp_max(sv);
            const float mn = fmaxf(m[g], mt);
            const float p = valid ? exp2f(sv - mn) : 0.f;
            alpha[g] = (m[g] == -INFINITY) ? 0.f : exp2f(m[g] - mn);
            l[g] = l[g] * alpha[g] + dsb::warp_sum(p);
            m[g] = mn;
            wp[g * PA_TILE + lane] = p;
        }
        __syncwarp();  // every lane is done with the K tile
#pragma unroll
        for (int i = 0; i < N_LOADS; i++) {
            const int r = i * ROWS_PER_LOAD + lane / LANES_PER_ROW, piece = lane % LANES_PER_ROW;
            *reinterpret_cast<uint4*>(wv + r * ROW + piece * 8) = vreg[i];
        }
        __syncwarp();
        // ---- O += P V: lane owns DV output columns ----
#pragma unroll
        for (int g = 0; g < G; g++)
#pragma unroll
            for (int c = 0; c < DV; c++) acc[g][c] *= alpha[g];
#pragma unroll 4
        for (int j = 0; j < PA_TILE; j += 4) {
            float4 pj[G];
#pragma unroll
            for (int g = 0; g < G; g++) pj[g] = *reinterpret_cast<const float4*>(wp + g * PA_TILE + j);
#pragma unroll
            for (int jj = 0; jj < 4; jj++) {
                float vf[DV];
                if constexpr (DV == 4) {
                    const uint2 raw = *reinterpret_cast<const uint2*>(wv + (j + jj) * ROW + lane * 4);
                    const __nv_bfloat162* h2 = reinterpret_cast<const __nv_bfloat162*>(&raw);
                    const float2 f0 = __bfloat1622float2(h2[0]), f1 = __bfloat1622float2(h2[1]);
                    vf[0] = f0.x, vf[1] = f0.y, vf[2] = f1.x, vf[3] = f1.y;
                } else {
                    const __nv_bfloat162 h2 = *reinterpret_cast<const __nv_bfloat162*>(wv + (j + jj) * ROW + lane * 2);
                    const float2 f0 = __bfloat1622float2(h2);
                    vf[0] = f0.x, vf[1] = f0.y;
                }
#pragma unroll
                for (int g = 0; g < G; g++) {
                    const float pv = jj == 0 ? pj[g].x : (jj == 1 ? pj[g].y : (jj == 2 ? pj[g].z : pj[g].w));
#pragma unroll
                    for (int c = 0; c < DV; c++) acc[g][c] = fmaf(pv, vf[c], acc[g][c]);
                }
            }
        }
        __syncwarp();
    }
    // ---- merge the warps of this CTA (the tile buffers now carry the partial outputs) ----
    float* my_acc = reinterpret_cast<float*>(s_kv[warp]);
#pragma unroll
    for (int g = 0; g < G; g++) {
        if (lane == 0) {
            s_m[warp][g] = m[g];
            s_l[warp][g] = l[g];
        }
#pragma unroll
        for (int c = 0; c < DV; c++) my_acc[g * D + lane * DV + c] = acc[g][c];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < G * D; i += PA_THREADS) {
        const int g = i / D, d = i % D;
        float mm = -INFINITY;
#pragma unroll
        for (int w = 0; w < PA_WARPS; w++) mm = fmaxf(mm, s_m[w][g]);
        float ll = 0.f, o = 0.f;
#pragma unroll
        for (int w = 0; w < PA_WARPS; w++) {
            const float f = (s_m[w][g] == -INFINITY) ? 0.f : exp2f(s_m[w][g] - mm);
            ll += s_l[w][g] * f;
            o += reinterpret_cast<const float*>(s_kv[w])[i] * f;
        }
        const int hq = hk * G + g;
        if (a.splits == 1) {
            a.out[((int64_t)t * a.Hq + hq) * D + d] = __float2bfloat16_rn(ll > 0.f ? o / ll : 0.f);
        } else {
            const int64_t base = ((int64_t)ti * a.Hq + hq) * a.splits + split;
            a.part_acc[base * D + d] = o;
            if (d == 0) {
                a.part_ml[base * 2] = mm;
                a.part_ml[base * 2 + 1] = ll;
            }
        }
    }
}

template <int D>
__global__ void __launch_bounds__(D) paged_attention_merge_kernel(const float* __restrict__ part_acc,
                                                                  const float* __restrict__ part_ml,
                                                                  __nv_bfloat16* __restrict__ out, int splits, int Hq,
                                                                  const int* __restrict__ tok_row)
{
    const int64_t row = blockIdx.x;  // listed token * Hq + head
    const int d = threadIdx.x;
    const int64_t out_row = tok_row != nullptr ? (int64_t)tok_row[row / Hq] * Hq + row % Hq : row;
    float mm = -INFINITY;
    for (int s = 0; s < splits; s++) mm = fmaxf(mm, part_ml[(row * splits + s) * 2]);
    float ll = 0.f, o = 0.f;
    for (int s = 0; s < splits; s++) {
        const float ms = part_ml[(row * splits + s) * 2];
        const float f = (ms == -INFINITY) ? 0.f : exp2f(ms - mm);
        ll += part_ml[(row * splits + s) * 2 + 1] * f;
        o += part_acc[(row * splits + s) * D + d] * f;
    }
    out[out_row * D + d] = __float2bfloat16_rn(ll > 0.f ? o / ll : 0.f);
}

// one CTA per token; threads sweep (head, rotary pair)
template <typename T>
__global__ void __launch_bounds__(256) kv_append_rope_kernel(T* __restrict__ q, const T* __restrict__ k,
                                                             const T* __restrict__ v, T* __restrict__ kv,
                                                             const float* __restrict__ cos_tab,
                                                             const float* __restrict__ sin_tab,
                                                             const int* __restrict__ tok_seq,
                                                             const int* __restrict__ tok_pos,
                                                             const int* __restrict__ block_table, int64_t q_stride,
                                                             int64_t k_stride, int64_t v_stride, int Hq, int Hkv, int D,
                                                             int rotary_dim, int block_size, int max_blocks)
{
    const int t = blockIdx.x;
    const int seq = tok_seq[t], pos = tok_pos[t];
    const int blk = block_table[(int64_t)seq * max_blocks + pos / block_size], off = pos % block_size;
    const int64_t plane = (int64_t)block_size * Hkv * D;
    T* kdst = kv + ((int64_t)blk * 2) * plane + (int64_t)off * Hkv * D;
    T* vdst = kdst + plane;
    const int half = rotary_dim / 2;
    const float* c = cos_tab + (int64_t)pos * half;
    const float* s = sin_tab + (int64_t)pos * half;
    // q heads: rotate in place
    for (int i = threadIdx.x; i < Hq * half; i += blockDim.x) {
        const int h = i / half, j = i % half;
        T* p = q + (int64_t)t * q_stride + (int64_t)h * D;
        const float x1 = dsb::to_f<T>(p[j]), x2 = dsb::to_f<T>(p[j + half]);
        p[j] = dsb::from_f<T>(x1 * c[j] - x2 * s[j]);
        p[j + half] = dsb::from_f<T>(x2 * c[j] + x1 * s[j]);
    }
    // k heads: rotate into the cache; pass-through dims and v: copy
    for (int i = threadIdx.x; i < Hkv * half; i += blockDim.x) {
        const int h = i / half, j = i % half;
        const T* p = k + (int64_t)t * k_stride + (int64_t)h * D;
        const float x1 = dsb::to_f<T>(p[j]), x2 = dsb::to_f<T>(p[j + half]);
        kdst[h * D + j] = dsb::from_f<T>(x1 * c[j] - x2 * s[j]);
        kdst[h * D + j + half] = dsb::from_f<T>(x2 * c[j] + x1 * s[j]);
    }
    const int rest = D - rotary_dim;
    for (int i = threadIdx.x; i < Hkv * rest; i += blockDim.x) {
        const int h = i / rest, j = rotary_dim + i % rest;
        kdst[h * D + j] = k[(int64_t)t * k_stride + (int64_t)h * D + j];
    }
    for (int i = threadIdx.x; i < Hkv * D; i += blockDim.x) vdst[i] = v[(int64_t)t * v_stride + i];
}

}  // namespace

extern "C" {

// q [T, Hq, D] (rotated in place), k / v [T, Hkv, D] (token strides in elements), kv: layer base of the blocked cache.
// rotary_dim == 0: no rotation (learned positions), k / v are only appended. dt: 1 fp16, 2 bf16.
int dsb200_ragged_kv_append_rope(void* q, const void* k, const void* v, void* kv, const float* cos_tab,
                                 const float* sin_tab, const int* tok_seq, const int* tok_pos, const int* block_table,
                                 int T, int64_t q_stride, int64_t k_stride, int64_t v_stride, int Hq, int Hkv, int D,
                                 int rotary_dim, int block_size, int max_blocks, int dt, cudaStream_t stream)
{
    if (T <= 0) return 0;
    if (rotary_dim % 2 != 0 || rotary_dim > D) {
        dsb200_set_error("ragged_kv_append_rope: rotary_dim must be even and <= head_dim");
        return -1;
    }
    if (dt == DSB_DT_BF16)
        kv_append_rope_kernel<__nv_bfloat16><<<T, 256, 0, stream>>>(
            (__nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v, (__nv_bfloat16*)kv, cos_tab, sin_tab,
            tok_seq, tok_pos, block_table, q_stride, k_stride, v_stride, Hq, Hkv, D, rotary_dim, block_size, max_blocks);
    else if (dt == DSB_DT_F16)
        kv_append_rope_kernel<__half><<<T, 256, 0, stream>>>((__half*)q, (const __half*)k, (const __half*)v, (__half*)kv,
                                                             cos_tab, sin_tab, tok_seq, tok_pos, block_table, q_stride,
                                                             k_stride, v_stride, Hq, Hkv, D, rotary_dim, block_size,
                                                             max_blocks);
    else {
        dsb200_set_error("ragged_kv_append_rope: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("ragged_kv_append_rope");
    return 0;
}

// The T listed tokens are rows tok_row[i] of q / out (tok_row null: row i), so a subset of a ragged batch can be served.
// Scratch needed for `splits` > 1: part_acc T*Hq*splits*D floats, part_ml T*Hq*splits*2 floats.
// Supported: bf16, head_dim 64 / 128, Hq / Hkv in {1, 2, 4, 8}. Returns the number of splits used (>= 1) or < 0.
int dsb200_ragged_paged_attention(const void* q, const void* kv, void* out, const int* tok_seq, const int* tok_pos,
                                  const int* block_table, const int* tok_row, float* part_acc, float* part_ml, int T,
                                  int64_t q_stride,
                                  int Hq, int Hkv, int D, int block_size, int max_blocks, int splits, float scale,
                                  cudaStream_t stream)
{
    if (T <= 0) return 1;
    const int G = Hkv > 0 ? Hq / Hkv : 0;
    if (Hkv <= 0 || Hq % Hkv != 0 || !(G == 1 || G == 2 || G == 4 || G == 8) || !(D == 64 || D == 128)) {
        dsb200_set_error("ragged_paged_attention: needs head_dim 64/128 and Hq/Hkv in {1,2,4,8}");
        return -1;
    }
    if (splits < 1) splits = 1;
    if (splits > 1 && (part_acc == nullptr || part_ml == nullptr)) {
        dsb200_set_error("ragged_paged_attention: split-KV needs the partial-result scratch buffers");
        return -1;
    }
    PagedArgs a;
    a.q = (const __nv_bfloat16*)q, a.kv = (const __nv_bfloat16*)kv, a.out = (__nv_bfloat16*)out;
    a.tok_seq = tok_seq, a.tok_pos = tok_pos, a.block_table = block_table, a.tok_row = tok_row;
    a.part_acc = part_acc, a.part_ml = part_ml;
    a.q_stride = q_stride;
    a.T = T, a.Hq = Hq, a.Hkv = Hkv, a.block_size = block_size, a.max_blocks = max_blocks, a.splits = splits;
    a.scale_log2 = scale * 1.4426950408889634f;
    dim3 grid(T, Hkv, splits);
#define PA_LAUNCH(DD, GG) paged_attention_kernel<DD, GG><<<grid, PA_THREADS, 0, stream>>>(a)
#define PA_SWITCH(DD)                 \
    do {                              \
        if (G == 1) PA_LAUNCH(DD, 1); \
        else if (G == 2) PA_LAUNCH(DD, 2); \
        else if (G == 4) PA_LAUNCH(DD, 4); \
        else PA_LAUNCH(DD, 8);        \
    } while (0)
    if (D == 128)
        PA_SWITCH(128);
    else
        PA_SWITCH(64);
#undef PA_SWITCH
#undef PA_LAUNCH
    DSB_CHECK_LAUNCH("ragged_paged_attention");
    if (splits > 1) {
        if (D == 128)
            paged_attention_merge_kernel<128><<<T * Hq, 128, 0, stream>>>(part_acc, part_ml, (__nv_bfloat16*)out, splits, Hq,
                                                                          tok_row);
        else
            paged_attention_merge_kernel<64><<<T * Hq, 64, 0, stream>>>(part_acc, part_ml, (__nv_bfloat16*)out, splits, Hq,
                                                                        tok_row);
        DSB_CHECK_LAUNCH("ragged_paged_attention_merge");
    }
    return splits;
}

}  // extern "C"
