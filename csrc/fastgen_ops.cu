// FastGen / inference helper kernels for sm_100a: embedding lookup for a ragged batch, last-token gather for the logits,
// bias + activation, gated activations, top-k expert gating, MoE scatter / weighted gather.
//
// Reference: inference/v2/kernels/ragged_ops/{embed,logits_gather,moe_scatter,moe_gather,top_k_gating}/*.cu,
// core_ops/{bias_activations,gated_activations}/*.cu; v1 csrc/transformer/inference/csrc/{gelu,relu}.cu
// (fused_bias_gelu / fused_bias_relu / fused_gate_activation).
//
// All of these are bandwidth-bound row operations: one CTA (or warp) owns a row, rows move as 16-byte vectors, math in
// fp32. dtype codes follow common.cuh (1 = fp16, 2 = bf16).
#include "common.cuh"

#include <type_traits>

namespace {

template <typename T>
struct Vec8 {
    uint4 raw;
    __device__ __forceinline__ float get(int i) const
    {
        const T* p = reinterpret_cast<const T*>(&raw);
        if constexpr (sizeof(T) == 2 && std::is_same<T, __nv_bfloat16>::value) return __bfloat162float(p[i]);
        else return __half2float(p[i]);
    }
    __device__ __forceinline__ void set(int i, float v)
    {
        T* p = reinterpret_cast<T*>(&raw);
        if constexpr (std::is_same<T, __nv_bfloat16>::value) p[i] = __float2bfloat16(v);
        else p[i] = __float2half(v);
    }
};

enum Act { ACT_IDENTITY = 0, ACT_RELU = 1, ACT_GELU = 2, ACT_GELU_TANH = 3, ACT_SILU = 4 };

__device__ __forceinline__ float activate(float x, int act)
{
    switch (act) {
        case ACT_RELU: return fmaxf(x, 0.f);
        case ACT_GELU: return 0.5f * x * (1.f + erff(x * 0.70710678118654752f));
        case ACT_GELU_TANH: {
            const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
            return 0.5f * x * (1.f + tanhf(u));
        }
        case ACT_SILU: return x / (1.f + __expf(-x));
        default: return x;
    }
}

// out[r, :] = table[ids[r], :] (+ pos_table[pos[r] + pos_offset, :]); cols % 8 == 0
template <typename T>
__global__ void __launch_bounds__(128) embed_rows_kernel(const T* __restrict__ table, const int64_t* __restrict__ ids,
                                                         const T* __restrict__ pos_table,
                                                         const int64_t* __restrict__ pos, int pos_offset,
                                                         T* __restrict__ out, int cols)
{
    const int64_t r = blockIdx.x;
    const uint4* src = reinterpret_cast<const uint4*>(table + ids[r] * (int64_t)cols);
    const uint4* psrc = pos_table ? reinterpret_cast<const uint4*>(pos_table + (pos[r] + pos_offset) * (int64_t)cols)
                                  : nullptr;
    uint4* dst = reinterpret_cast<uint4*>(out + r * (int64_t)cols);
    for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) {
        Vec8<T> v;
        v.raw = __ldg(src + c);
        if (psrc) {
            Vec8<T> p;
            p.raw = __ldg(psrc + c);
#pragma unroll
            for (int i = 0; i < 8; i++) v.set(i, v.get(i) + p.get(i));
        }
        dst[c] = v.raw;
    }
}

// y = act(x + bias): x [rows, cols]
template <typename T>
__global__ void __launch_bounds__(256) bias_act_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                       T* __restrict__ y, int64_t total_vec, int cols, int act)
{
    const int cv = cols / 8;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total_vec; i += (int64_t)gridDim.x * blockDim.x) {
        Vec8<T> v, b;
        v.raw = reinterpret_cast<const uint4*>(x)[i];
        if (bias) b.raw = __ldg(reinterpret_cast<const uint4*>(bias) + (i % cv));
#pragma unroll
        for (int k = 0; k < 8; k++) v.set(k, activate(v.get(k) + (bias ? b.get(k) : 0.f), act));
        reinterpret_cast<uint4*>(y)[i] = v.raw;
    }
}

// y[r, :] = act(x[r, :inter] + b[:inter]) * (x[r, inter:] + b[inter:]); x [rows, 2 * inter]
template <typename T>
__global__ void __launch_bounds__(256) gated_act_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                        T* __restrict__ y, int64_t rows, int inter, int act)
{
    const int iv = inter / 8;
    const int64_t total = rows * iv;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / iv;
        const int c = (int)(i % iv);
        const uint4* row = reinterpret_cast<const uint4*>(x + r * 2 * (int64_t)inter);
        Vec8<T> g, u, bg, bu, o;
        g.raw = row[c];
        u.raw = row[iv + c];
        if (bias) {
            bg.raw = __ldg(reinterpret_cast<const uint4*>(bias) + c);
            bu.raw = __ldg(reinterpret_cast<const uint4*>(bias) + iv + c);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
            const float gv = g.get(k) + (bias ? bg.get(k) : 0.f), uv = u.get(k) + (bias ? bu.get(k) : 0.f);
            o.set(k, activate(gv, act) * uv);
        }
        reinterpret_cast<uint4*>(y + r * (int64_t)inter)[c] = o.raw;
    }
}

// One warp per token: softmax over E logits (fp32), the k largest, optional renormalisation, per-expert counts.
// E <= 1024, k <= 16. assignment[t, j] = position of (t, j) among the tokens of its expert (atomic counter order).
__global__ void __launch_bounds__(128) topk_gating_kernel(const float* __restrict__ logits, int T, int E, int k,
                                                          int renorm, float* __restrict__ scores,
                                                          int* __restrict__ experts, int* __restrict__ counts,
                                                          int* __restrict__ slot)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= T) return;
    const float* row = logits + (int64_t)warp * E;
    float mx = -INFINITY;
    for (int e = lane; e < E; e += 32) mx = fmaxf(mx, row[e]);
    mx = dsb::warp_max(mx);
    float sum = 0.f;
    for (int e = lane; e < E; e += 32) sum += __expf(row[e] - mx);
    sum = dsb::warp_sum(sum);
    const float inv = 1.f / sum;
    unsigned long long taken_lo = 0;  // per-lane mask over this lane's (up to 32) strided entries
    float picked[16];
    int picked_e[16];
    float psum = 0.f;
    for (int j = 0; j < k; j++) {
        float best = -INFINITY;
        int best_e = 0x7fffffff;
        for (int e = lane, i = 0; e < E; e += 32, i++) {
            if (taken_lo >> i & 1ull) continue;
            const float v = row[e];
            if (v > best || (v == best && e < best_e)) { best = v; best_e = e; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oe = __shfl_xor_sync(0xffffffffu, best_e, o);
            if (ob > best || (ob == best && oe < best_e)) { best = ob; best_e = oe; }
        }
        if ((best_e & 31) == lane) taken_lo |= 1ull << (best_e >> 5);
        picked[j] = __expf(best - mx) * inv;
        picked_e[j] = best_e;
        psum += picked[j];
    }
    if (lane == 0) {
        const float scale = renorm ? 1.f / fmaxf(psum, 1e-20f) : 1.f;
        for (int j = 0; j < k; j++) {
            scores[(int64_t)warp * k + j] = picked[j] * scale;
            experts[(int64_t)warp * k + j] = picked_e[j];
            const int s = atomicAdd(counts + picked_e[j], 1);
            if (slot) slot[(int64_t)warp * k + j] = s;
        }
    }
}

// out[dst[t, j], :] = x[t, :] for every (t, j) with dst >= 0
template <typename T>
__global__ void __launch_bounds__(128) moe_scatter_kernel(const T* __restrict__ x, const int* __restrict__ dst, int k,
                                                          T* __restrict__ out, int cols)
{
    const int64_t t = blockIdx.x;
    const uint4* src = reinterpret_cast<const uint4*>(x + t * (int64_t)cols);
    for (int j = 0; j < k; j++) {
        const int d = dst[t * k + j];
        if (d < 0) continue;
        uint4* o = reinterpret_cast<uint4*>(out + (int64_t)d * cols);
        for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) o[c] = __ldg(src + c);
    }
}

// out[t, :] = sum_j w[t, j] * expert_out[src[t, j], :]
template <typename T>
__global__ void __launch_bounds__(128) moe_gather_kernel(const T* __restrict__ expert_out, const int* __restrict__ src,
                                                         const float* __restrict__ w, int k, T* __restrict__ out,
                                                         int cols)
{
    const int64_t t = blockIdx.x;
    for (int c = threadIdx.x; c < cols / 8; c += blockDim.x) {
        float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int j = 0; j < k; j++) {
            const int s = src[t * k + j];
            if (s < 0) continue;
            const float wj = w[t * k + j];
            Vec8<T> v;
            v.raw = __ldg(reinterpret_cast<const uint4*>(expert_out + (int64_t)s * cols) + c);
#pragma unroll
            for (int i = 0; i < 8; i++) acc[i] += wj * v.get(i);
        }
        Vec8<T> o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.set(i, acc[i]);
        reinterpret_cast<uint4*>(out + t * (int64_t)cols)[c] = o.raw;
    }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace

namespace {
template <typename T>
int launch_embed(const void* table, const int64_t* ids, const void* pos_table, const int64_t* pos, int pos_offset,
                 void* out, int64_t rows, int cols, cudaStream_t stream)
{
    embed_rows_kernel<T><<<(unsigned)rows, 128, 0, stream>>>((const T*)table, ids, (const T*)pos_table, pos, pos_offset,
                                                             (T*)out, cols);
    return 0;
}

inline unsigned grid_for(int64_t work, int threads)
{
    const int64_t blocks = (work + threads - 1) / threads;
    const int64_t cap = (int64_t)DSB_NUM_SMS * 8;
    return (unsigned)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}
}  // namespace

extern "C" {

// out[r] = table[ids[r]] (+ pos_table[pos[r] + pos_offset]). Also the logits gather (table = hidden states, ids = last
// token of every sequence). ids / pos are int64 on the device.
int dsb200_embed_rows(const void* table, const int64_t* ids, const void* pos_table, const int64_t* pos, int pos_offset,
                      void* out, int64_t rows, int cols, int dt, cudaStream_t stream)
{
    if (rows == 0) return 0;
    if (cols % 8 || !aligned16(table) || !aligned16(out) || (pos_table && !aligned16(pos_table))) {
        dsb200_set_error("embed_rows: rows must be 16-byte aligned and a multiple of 8 elements wide");
        return -1;
    }
    if (dt == DSB_DT_BF16) launch_embed<__nv_bfloat16>(table, ids, pos_table, pos, pos_offset, out, rows, cols, stream);
    else if (dt == DSB_DT_F16) launch_embed<__half>(table, ids, pos_table, pos, pos_offset, out, rows, cols, stream);
    else {
        dsb200_set_error("embed_rows: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("embed_rows");
    return 0;
}

// y = act(x + bias); act: 0 identity, 1 relu, 2 gelu (erf), 3 gelu (tanh), 4 silu. bias may be null.
int dsb200_bias_act(const void* x, const void* bias, void* y, int64_t rows, int cols, int act, int dt,
                    cudaStream_t stream)
{
    if (rows == 0) return 0;
    if (cols % 8 || !aligned16(x) || !aligned16(y) || (bias && !aligned16(bias))) {
        dsb200_set_error("bias_act: tensors must be 16-byte aligned and a multiple of 8 elements wide");
        return -1;
    }
    const int64_t total = rows * (cols / 8);
    const unsigned grid = grid_for(total, 256);
    if (dt == DSB_DT_BF16)
        bias_act_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)bias,
                                                                 (__nv_bfloat16*)y, total, cols, act);
    else if (dt == DSB_DT_F16)
        bias_act_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)x, (const __half*)bias, (__half*)y, total, cols,
                                                          act);
    else {
        dsb200_set_error("bias_act: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("bias_act");
    return 0;
}

// y [rows, inter] = act(x[:, :inter] + b[:inter]) * (x[:, inter:] + b[inter:]); x [rows, 2 * inter] contiguous.
int dsb200_gated_act(const void* x, const void* bias, void* y, int64_t rows, int inter, int act, int dt,
                     cudaStream_t stream)
{
    if (rows == 0) return 0;
    if (inter % 8 || !aligned16(x) || !aligned16(y) || (bias && !aligned16(bias))) {
        dsb200_set_error("gated_act: tensors must be 16-byte aligned, intermediate size a multiple of 8");
        return -1;
    }
    const unsigned grid = grid_for(rows * (inter / 8), 256);
    if (dt == DSB_DT_BF16)
        gated_act_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)bias,
                                                                  (__nv_bfloat16*)y, rows, inter, act);
    else if (dt == DSB_DT_F16)
        gated_act_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)x, (const __half*)bias, (__half*)y, rows, inter,
                                                           act);
    else {
        dsb200_set_error("gated_act: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("gated_act");
    return 0;
}

// Softmax + top-k per token. logits fp32 [T, E]; scores fp32 [T, k]; experts int32 [T, k]; counts int32 [E] (must be
// zeroed by the caller); slot int32 [T, k] or null: arrival order of the assignment inside its expert.
int dsb200_topk_gating(const float* logits, int T, int E, int k, int renorm, float* scores, int* experts, int* counts,
                       int* slot, cudaStream_t stream)
{
    if (T == 0) return 0;
    if (E < 1 || E > 1024 || k < 1 || k > 16 || k > E) {
        dsb200_set_error("topk_gating: need 1 <= k <= min(16, E) and E <= 1024");
        return -1;
    }
    const int warps_per_block = 4;
    topk_gating_kernel<<<(T + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, stream>>>(
        logits, T, E, k, renorm, scores, experts, counts, slot);
    DSB_CHECK_LAUNCH("topk_gating");
    return 0;
}

// out[dst[t, j]] = x[t] for dst >= 0. x [T, cols], dst int32 [T, k], out [R, cols].
int dsb200_moe_scatter(const void* x, const int* dst, int T, int k, void* out, int cols, int dt, cudaStream_t stream)
{
    if (T == 0) return 0;
    if (cols % 8 || !aligned16(x) || !aligned16(out)) {
        dsb200_set_error("moe_scatter: rows must be 16-byte aligned and a multiple of 8 elements wide");
        return -1;
    }
    if (dt == DSB_DT_BF16)
        moe_scatter_kernel<__nv_bfloat16><<<T, 128, 0, stream>>>((const __nv_bfloat16*)x, dst, k, (__nv_bfloat16*)out, cols);
    else if (dt == DSB_DT_F16)
        moe_scatter_kernel<__half><<<T, 128, 0, stream>>>((const __half*)x, dst, k, (__half*)out, cols);
    else {
        dsb200_set_error("moe_scatter: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("moe_scatter");
    return 0;
}

// out[t] = sum_j w[t, j] * expert_out[src[t, j]] (src < 0: dropped). w fp32 [T, k].
int dsb200_moe_gather(const void* expert_out, const int* src, const float* w, int T, int k, void* out, int cols, int dt,
                      cudaStream_t stream)
{
    if (T == 0) return 0;
    if (cols % 8 || !aligned16(expert_out) || !aligned16(out)) {
        dsb200_set_error("moe_gather: rows must be 16-byte aligned and a multiple of 8 elements wide");
        return -1;
    }
    if (dt == DSB_DT_BF16)
        moe_gather_kernel<__nv_bfloat16><<<T, 128, 0, stream>>>((const __nv_bfloat16*)expert_out, src, w, k,
                                                                (__nv_bfloat16*)out, cols);
    else if (dt == DSB_DT_F16)
        moe_gather_kernel<__half><<<T, 128, 0, stream>>>((const __half*)expert_out, src, w, k, (__half*)out, cols);
    else {
        dsb200_set_error("moe_gather: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("moe_gather");
    return 0;
}

}  // extern "C"
