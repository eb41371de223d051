// RoPE, SwiGLU and fused softmax-cross-entropy for sm_100a (all HBM-bound, 16-byte accesses, fp32 math).
//
// Reference counterparts: apply_rotary_pos_half (csrc/transformer/inference/csrc/apply_rotary_pos_emb.cu:23-169,
// inference only), Triton _swiglu_fwd/_bwd (ops/triton_ops/swiglu_triton.py:36-70), TiledFusedLogitsLoss
// (runtime/sequence_parallel/ulysses_sp.py:1065, eager PyTorch per tile). Here they are training kernels with
// backward, and the loss kernel writes d(logits) in place so the fp32 [tokens, vocab] tensor never exists.
#include "common.cuh"

namespace {

// ---------------------------------------------------------------------------------------------------------
// RoPE (half-rotation, HF Llama convention): for j < D/2:
//   out[j]       = x[j]*cos[j] - x[j+D/2]*sin[j]
//   out[j+D/2]   = x[j+D/2]*cos[j] + x[j]*sin[j]
// backward == forward with sin -> -sin. x: [tokens, heads, D] (row stride given), cos/sin: fp32 [positions, D/2],
// pos: int32 [tokens] or null (then position = token % seq_len). In-place allowed.
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(T* __restrict__ x, const float* __restrict__ cs,
                                                   const float* __restrict__ sn, const int* __restrict__ pos,
                                                   int tokens, int heads, int dim, int64_t token_stride, int seq_len,
                                                   float sign)
{
    const int half = dim >> 1;
    const int vec_per_head = half >> 3;  // 8 elements per thread from each half
    const int64_t total = (int64_t)tokens * heads * vec_per_head;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int v = idx % vec_per_head;
        const int h = (idx / vec_per_head) % heads;
        const int t = idx / ((int64_t)vec_per_head * heads);
        const int p = pos != nullptr ? pos[t] : (t % seq_len);
        T* base = x + (int64_t)t * token_stride + (int64_t)h * dim + v * 8;
        dsb::f8 a = dsb::load8(base), b = dsb::load8(base + half);
        const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)p * half + v * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)p * half + v * 8);
        float4 c0 = c4[0], c1 = c4[1], s0 = s4[0], s1 = s4[1];
        const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
        const float s[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
        dsb::f8 oa, ob;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float si = s[i] * sign;
            oa.v[i] = a.v[i] * c[i] - b.v[i] * si;
            ob.v[i] = b.v[i] * c[i] + a.v[i] * si;
        }
        dsb::store8(base, oa);
        dsb::store8(base + half, ob);
    }
}

// ---------------------------------------------------------------------------------------------------------
// SwiGLU: y = silu(gate) * up. Inputs may be the two halves of one fused [rows, 2*I] tensor (stride = 2*I).
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + __expf(-x)); }

template <typename T>
__global__ void __launch_bounds__(256) swiglu_fwd_kernel(const T* __restrict__ gate, const T* __restrict__ up,
                                                         T* __restrict__ y, int64_t rows, int inter, int64_t in_stride)
{
    const int vpr = inter >> 3;
    const int64_t total = rows * vpr;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = idx / vpr;
        const int v = idx % vpr;
        dsb::f8 g = dsb::load8(gate + r * in_stride + v * 8), u = dsb::load8(up + r * in_stride + v * 8), o;
#pragma unroll
        for (int i = 0; i < 8; i++) o.v[i] = g.v[i] * sigmoidf_(g.v[i]) * u.v[i];
        dsb::store8(y + r * inter + v * 8, o);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) swiglu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ gate,
                                                         const T* __restrict__ up, T* __restrict__ dgate,
                                                         T* __restrict__ dup, int64_t rows, int inter,
                                                         int64_t in_stride)
{
    const int vpr = inter >> 3;
    const int64_t total = rows * vpr;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total;
         idx += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = idx / vpr;
        const int v = idx % vpr;
        dsb::f8 d = dsb::load8(dy + r * inter + v * 8);
        dsb::f8 g = dsb::load8(gate + r * in_stride + v * 8), u = dsb::load8(up + r * in_stride + v * 8), og, ou;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const float s = sigmoidf_(g.v[i]);
            const float silu = g.v[i] * s;
            ou.v[i] = d.v[i] * silu;
            og.v[i] = d.v[i] * u.v[i] * (s + silu * (1.f - s));
        }
        dsb::store8(dgate + r * in_stride + v * 8, og);
        dsb::store8(dup + r * in_stride + v * 8, ou);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Softmax cross-entropy over the vocabulary, one CTA per token row.
//   pass 1: online max / sum-exp (one read of the row)     -> loss[row] = lse - logit[target], lse[row]
//   pass 2 (optional, `write_grad`): logits <- (softmax - onehot) * grad_scale   (row is L2-resident: 256 KB)
// target == ignore_index -> loss 0, gradient 0.
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) ce_fwd_kernel(T* __restrict__ logits, const int64_t* __restrict__ target,
                                                      float* __restrict__ loss, float* __restrict__ lse_out, int rows,
                                                      int vocab, int64_t row_stride, int ignore_index, int write_grad,
                                                      float grad_scale, const float* __restrict__ grad_scale_dev)
{
    __shared__ float red[33];
    if (grad_scale_dev != nullptr) grad_scale *= *grad_scale_dev;  // e.g. 1 / (number of valid tokens), kept on device
    const int row = blockIdx.x;
    T* lr = logits + (int64_t)row * row_stride;
    const int64_t tgt = target[row];
    const int nvec = vocab >> 3;
    float m = -INFINITY, s = 0.f;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
        dsb::f8 a = dsb::load8(lr + v * 8);
        float lm = a.v[0];
#pragma unroll
        for (int i = 1; i < 8; i++) lm = fmaxf(lm, a.v[i]);
        const float nm = fmaxf(m, lm);
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 8; i++) acc += __expf(a.v[i] - nm);
        s = s * __expf(m - nm) + acc;
        m = nm;
    }
    for (int i = (nvec << 3) + threadIdx.x; i < vocab; i += blockDim.x) {
        const float a = dsb::to_f<T>(lr[i]);
        const float nm = fmaxf(m, a);
        s = s * __expf(m - nm) + __expf(a - nm);
        m = nm;
    }
    const float gm = dsb::block_max(m, red);
    const float part = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = dsb::block_sum(part, red);
    const float lse = gm + __logf(gs);
    const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= vocab;
    if (threadIdx.x == 0) {
        const float lt = ignored ? 0.f : dsb::to_f<T>(lr[tgt]);
        loss[row] = ignored ? 0.f : (lse - lt);
        if (lse_out != nullptr) lse_out[row] = lse;
    }
    if (!write_grad) return;
    __syncthreads();  // the target logit has been read
    const float gsc = ignored ? 0.f : grad_scale;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x) {
        dsb::f8 a = dsb::load8(lr + v * 8), o;
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int col = v * 8 + i;
            o.v[i] = (__expf(a.v[i] - lse) - ((col == tgt) ? 1.f : 0.f)) * gsc;
        }
        dsb::store8(lr + v * 8, o);
    }
    for (int i = (nvec << 3) + threadIdx.x; i < vocab; i += blockDim.x) {
        const float a = dsb::to_f<T>(lr[i]);
        lr[i] = dsb::from_f<T>((__expf(a - lse) - ((i == tgt) ? 1.f : 0.f)) * gsc);
    }
}

// x[i] *= *scale (device scalar), vectorised; nothing is touched when *scale == 1
template <typename T>
__global__ void __launch_bounds__(256) scale_by_dev_kernel(T* __restrict__ x, int64_t n, const float* __restrict__ scale)
{
    const float s = *scale;
    if (s == 1.0f) return;
    const int64_t nvec = n >> 3;
    for (int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * blockDim.x) {
        dsb::f8 a = dsb::load8(x + v * 8);
#pragma unroll
        for (int i = 0; i < 8; i++) a.v[i] *= s;
        dsb::store8(x + v * 8, a);
    }
    if (blockIdx.x == 0)
        for (int64_t i = (nvec << 3) + threadIdx.x; i < n; i += blockDim.x) x[i] = dsb::from_f<T>(dsb::to_f<T>(x[i]) * s);
}

int grid_for(int64_t work_items, int threads, int per_sm)
{
    int64_t g = (work_items + threads - 1) / threads;
    int64_t cap = (int64_t)DSB_NUM_SMS * per_sm;
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace

extern "C" {

int dsb200_rope(void* x, const float* cos_tab, const float* sin_tab, const int* pos, int tokens, int heads, int dim,
                int64_t token_stride, int seq_len, int backward, int dt, cudaStream_t stream)
{
    if (tokens <= 0) return 0;
    if (dim % 16 != 0) {
        dsb200_set_error("rope: head_dim must be a multiple of 16");
        return -1;
    }
    const int64_t total = (int64_t)tokens * heads * (dim / 16);
    const int grid = grid_for(total, 256, 8);
    const float sign = backward ? -1.f : 1.f;
    if (dt == DSB_DT_BF16)
        rope_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16*)x, cos_tab, sin_tab, pos, tokens, heads, dim,
                                                             token_stride, seq_len, sign);
    else if (dt == DSB_DT_F16)
        rope_kernel<__half><<<grid, 256, 0, stream>>>((__half*)x, cos_tab, sin_tab, pos, tokens, heads, dim,
                                                      token_stride, seq_len, sign);
    else {
        dsb200_set_error("rope: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("rope");
    return 0;
}

int dsb200_swiglu_fwd(const void* gate, const void* up, void* y, int64_t rows, int inter, int64_t in_stride, int dt,
                      cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (inter % 8 != 0) {
        dsb200_set_error("swiglu: intermediate size must be a multiple of 8");
        return -1;
    }
    const int grid = grid_for(rows * (inter / 8), 256, 8);
    if (dt == DSB_DT_BF16)
        swiglu_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)gate, (const __nv_bfloat16*)up,
                                                                   (__nv_bfloat16*)y, rows, inter, in_stride);
    else if (dt == DSB_DT_F16)
        swiglu_fwd_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)gate, (const __half*)up, (__half*)y, rows,
                                                            inter, in_stride);
    else {
        dsb200_set_error("swiglu: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("swiglu_fwd");
    return 0;
}

int dsb200_swiglu_bwd(const void* dy, const void* gate, const void* up, void* dgate, void* dup, int64_t rows, int inter,
                      int64_t in_stride, int dt, cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (inter % 8 != 0) {
        dsb200_set_error("swiglu: intermediate size must be a multiple of 8");
        return -1;
    }
    const int grid = grid_for(rows * (inter / 8), 256, 8);
    if (dt == DSB_DT_BF16)
        swiglu_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(
            (const __nv_bfloat16*)dy, (const __nv_bfloat16*)gate, (const __nv_bfloat16*)up, (__nv_bfloat16*)dgate,
            (__nv_bfloat16*)dup, rows, inter, in_stride);
    else if (dt == DSB_DT_F16)
        swiglu_bwd_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)dy, (const __half*)gate, (const __half*)up,
                                                            (__half*)dgate, (__half*)dup, rows, inter, in_stride);
    else {
        dsb200_set_error("swiglu: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("swiglu_bwd");
    return 0;
}

int dsb200_ce_fwd_ex(void* logits, const int64_t* target, float* loss, float* lse, int rows, int vocab,
                     int64_t row_stride, int ignore_index, int write_grad, float grad_scale, const float* grad_scale_dev,
                     int dt, cudaStream_t stream);

// logits: [rows, vocab] with row stride `row_stride` elements (16-byte aligned rows); target int64 [rows].
int dsb200_ce_fwd(void* logits, const int64_t* target, float* loss, float* lse, int rows, int vocab, int64_t row_stride,
                  int ignore_index, int write_grad, float grad_scale, int dt, cudaStream_t stream)
{
    return dsb200_ce_fwd_ex(logits, target, loss, lse, rows, vocab, row_stride, ignore_index, write_grad, grad_scale,
                            nullptr, dt, stream);
}

// x (fp16 / bf16, 16-byte aligned, contiguous) *= *scale_dev; a no-op kernel when the device scalar is 1.
int dsb200_scale_by_dev(void* x, int64_t n, int dt, const float* scale_dev, cudaStream_t stream)
{
    if (n <= 0) return 0;
    const int grid = grid_for(n / 8 + 1, 256, 8);
    if (dt == DSB_DT_BF16)
        scale_by_dev_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16*)x, n, scale_dev);
    else if (dt == DSB_DT_F16)
        scale_by_dev_kernel<__half><<<grid, 256, 0, stream>>>((__half*)x, n, scale_dev);
    else {
        dsb200_set_error("scale_by_dev: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("scale_by_dev");
    return 0;
}

// as dsb200_ce_fwd; the gradient scale is grad_scale * (*grad_scale_dev) when the device pointer is given
int dsb200_ce_fwd_ex(void* logits, const int64_t* target, float* loss, float* lse, int rows, int vocab,
                     int64_t row_stride, int ignore_index, int write_grad, float grad_scale, const float* grad_scale_dev,
                     int dt, cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (row_stride % 8 != 0) {
        dsb200_set_error("cross_entropy: row stride must be a multiple of 8 elements");
        return -1;
    }
    const int threads = vocab >= 16384 ? 1024 : (vocab >= 2048 ? 256 : 128);
    if (dt == DSB_DT_BF16)
        ce_fwd_kernel<__nv_bfloat16><<<rows, threads, 0, stream>>>((__nv_bfloat16*)logits, target, loss, lse, rows, vocab,
                                                                   row_stride, ignore_index, write_grad, grad_scale,
                                                                   grad_scale_dev);
    else if (dt == DSB_DT_F16)
        ce_fwd_kernel<__half><<<rows, threads, 0, stream>>>((__half*)logits, target, loss, lse, rows, vocab, row_stride,
                                                            ignore_index, write_grad, grad_scale, grad_scale_dev);
    else {
        dsb200_set_error("cross_entropy: dtype must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("ce_fwd");
    return 0;
}

}  // extern "C"
