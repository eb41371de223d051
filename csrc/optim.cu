// Multi-tensor optimizers (Adam/AdamW, Lion, Adagrad, SGD-momentum), LAMB, and fp32 sum-of-squares reductions.
//
// What the reference does: csrc/adam/multi_tensor_adam.cu:33-127 + multi_tensor_apply.cuh:38-125 pack <=36
// tensor pointers / <=320 blocks into kernel *arguments* per launch, so a model with N tensors needs N/36
// launches and the pointer lists are rebuilt on the host every step; unscale, clip, fp32->bf16 copy are separate
// passes (stage_1_and_2.py:2350-2386).
// B200 design: one device-resident chunk table (built once while pointers are stable) -> ONE launch per step for
// every tensor; the kernel fuses grad unscale*clip, the optimizer math in fp32, and the low-precision weight
// write-back. Purely HBM-bound: every operand is touched once with 16-byte accesses
// (fp32: p,m,v read+write, grad read, bf16 write ~= 30 B/element).
#include "common.cuh"

namespace {

struct __align__(8) Chunk {
    void* p;
    const void* g;
    void* m;
    void* v;
    void* lp;
    int n;
    int group;
    int dt;  // p | g<<4 | state<<8 | lp<<12
    int pad;
};

constexpr int MAX_GROUPS = 64;
struct Hyper {
    float h[MAX_GROUPS][8];  // lr, beta1, beta2, eps, wd, bc1, bc2, mode
};

enum { K_ADAM = 0, K_LION = 1, K_ADAGRAD = 2, K_SGD = 3 };

template <int KIND>
__device__ __forceinline__ void update1(float& p, float g, float& m, float& v, const float* h)
{
    const float lr = h[0], b1 = h[1], b2 = h[2], eps = h[3], wd = h[4], bc1 = h[5], bc2 = h[6];
    const int mode = (int)h[7];
    if (KIND == K_ADAM) {
        if (mode == 0) g = fmaf(wd, p, g);  // L2 regularisation
        m = fmaf(b1, m, (1.f - b1) * g);
        v = fmaf(b2, v, (1.f - b2) * g * g);
        float denom = sqrtf(v / bc2) + eps;
        float upd = (m / bc1) / denom;
        if (mode == 1) upd = fmaf(wd, p, upd);  // decoupled (AdamW)
        p = fmaf(-lr, upd, p);
    } else if (KIND == K_LION) {
        p = p * (1.f - lr * wd);
        float c = fmaf(b1, m, (1.f - b1) * g);
        float s = (c > 0.f) ? 1.f : ((c < 0.f) ? -1.f : 0.f);
        p = fmaf(-lr, s, p);
        m = fmaf(b2, m, (1.f - b2) * g);
    } else if (KIND == K_ADAGRAD) {
        g = fmaf(wd, p, g);
        v = fmaf(g, g, v);
        p = fmaf(-lr, g / (sqrtf(v) + eps), p);
    } else {  // SGD with momentum in m; b2==1 -> nesterov; mode==1 -> first step (buffer = grad)
        g = fmaf(wd, p, g);
        if (b1 != 0.f) {
            m = (mode == 1) ? g : fmaf(b1, m, g);
            g = (b2 == 1.f) ? fmaf(b1, m, g) : m;
        }
        p = fmaf(-lr, g, p);
    }
}

template <int KIND>
__global__ void __launch_bounds__(512) multi_tensor_optim_kernel(const Chunk* __restrict__ table, Hyper hyper,
                                                                 float grad_scale)
{
    const Chunk c = table[blockIdx.x];
    const float* h = hyper.h[c.group];
    const int pdt = c.dt & 15, gdt = (c.dt >> 4) & 15, sdt = (c.dt >> 8) & 15, ldt = (c.dt >> 12) & 15;
    constexpr bool HAS_M = (KIND != K_ADAGRAD);
    constexpr bool HAS_V = (KIND == K_ADAM || KIND == K_ADAGRAD);
    const bool vec_ok = dsb::aligned16(c.p) && dsb::aligned16(c.g) && dsb::aligned16(c.m) &&
                        (!HAS_V || dsb::aligned16(c.v)) && (c.lp == nullptr || dsb::aligned16(c.lp));
    const int n4 = vec_ok ? (c.n >> 2) : 0;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 p = dsb::ld4_dt(c.p, pdt, i);
        float4 g = dsb::ld4_dt(c.g, gdt, i);
        float4 m = HAS_M ? dsb::ld4_dt(c.m, sdt, i) : make_float4(0, 0, 0, 0);
        float4 v = HAS_V ? dsb::ld4_dt(c.v, sdt, i) : make_float4(0, 0, 0, 0);
        g.x *= grad_scale; g.y *= grad_scale; g.z *= grad_scale; g.w *= grad_scale;
        update1<KIND>(p.x, g.x, m.x, v.x, h);
        update1<KIND>(p.y, g.y, m.y, v.y, h);
        update1<KIND>(p.z, g.z, m.z, v.z, h);
        update1<KIND>(p.w, g.w, m.w, v.w, h);
        dsb::st4_dt(c.p, pdt, i, p);
        if (HAS_M) dsb::st4_dt(c.m, sdt, i, m);
        if (HAS_V) dsb::st4_dt(c.v, sdt, i, v);
        if (c.lp != nullptr) dsb::st4_dt(c.lp, ldt, i, p);
    }
    for (int i = (n4 << 2) + threadIdx.x; i < c.n; i += blockDim.x) {
        float p = dsb::ld_dt(c.p, pdt, i), g = dsb::ld_dt(c.g, gdt, i) * grad_scale;
        float m = HAS_M ? dsb::ld_dt(c.m, sdt, i) : 0.f, v = HAS_V ? dsb::ld_dt(c.v, sdt, i) : 0.f;
        update1<KIND>(p, g, m, v, h);
        dsb::st_dt(c.p, pdt, i, p);
        if (HAS_M) dsb::st_dt(c.m, sdt, i, m);
        if (HAS_V) dsb::st_dt(c.v, sdt, i, v);
        if (c.lp != nullptr) dsb::st_dt(c.lp, ldt, i, p);
    }
}

// ---------------------------------------------------------------------------------------------------------
// sum of squares: stage 1 writes one partial per block into `scratch`, the last block to finish (ticket in
// scratch[SUMSQ_MAX_BLOCKS]) reduces the partials in fixed order -> deterministic result, one launch.
// ---------------------------------------------------------------------------------------------------------
constexpr int SUMSQ_MAX_BLOCKS = 1024;

template <typename T>
__device__ __forceinline__ float sumsq_body(const T* __restrict__ x, int64_t n)
{
    float acc = 0.f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    constexpr int VEC = 16 / sizeof(T);
    if (dsb::aligned16(x)) {
        const int64_t nv = n / VEC;
        const uint4* xv = reinterpret_cast<const uint4*>(x);
        for (int64_t i = tid; i < nv; i += nth) {
            uint4 raw = dsb::ldg_stream(xv + i);
            const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
            for (int k = 0; k < VEC; k++) {
                float f = dsb::to_f<T>(e[k]);
                acc = fmaf(f, f, acc);
            }
        }
        for (int64_t i = nv * VEC + tid; i < n; i += nth) {
            float f = dsb::to_f<T>(x[i]);
            acc = fmaf(f, f, acc);
        }
    } else {
        for (int64_t i = tid; i < n; i += nth) {
            float f = dsb::to_f<T>(x[i]);
            acc = fmaf(f, f, acc);
        }
    }
    return acc;
}

template <typename T>
__global__ void __launch_bounds__(512) sumsq_kernel(const T* __restrict__ x, int64_t n, float* __restrict__ scratch,
                                                    float* __restrict__ out)
{
    __shared__ float red[33];
    __shared__ bool is_last;
    float acc = sumsq_body<T>(x, n);
    acc = dsb::block_sum(acc, red);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + SUMSQ_MAX_BLOCKS);
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = acc;
        __threadfence();
        unsigned int t = atomicInc(ticket, gridDim.x - 1);  // wraps to 0 after the last block -> self-resetting
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) s += __ldcg(scratch + i);
        s = dsb::block_sum(s, red);
        if (threadIdx.x == 0) *out = s;
    }
}

// ---------------------------------------------------------------------------------------------------------
// LAMB: phase 1 computes the Adam-style update into `upd` (reusing the grad buffer) + per-block partial norms,
// phase 2 reduces the partials to the trust ratio and applies it. Reference: fused_lamb_cuda_kernel.cu
// lamb_cuda_kernel_part1/2/3 (:357-453).
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512) lamb_phase1(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t n, float* __restrict__ partials,
                                                   float b1, float b2, float eps, float wd, float bc1, float bc2,
                                                   int eps_mode)
{
    __shared__ float red[33];
    float wn = 0.f, un = 0.f;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = tid; i < n; i += nth) {
        float pi = p[i], gi = g[i];
        float mi = fmaf(b1, m[i], (1.f - b1) * gi);
        float vi = fmaf(b2, v[i], (1.f - b2) * gi * gi);
        m[i] = mi;
        v[i] = vi;
        float mh = mi / bc1, vh = vi / bc2;
        float denom = (eps_mode == 0) ? sqrtf(vh + eps) : (sqrtf(vh) + eps);
        float u = mh / denom + wd * pi;
        g[i] = u;  // stash the update in the (consumed) grad buffer
        wn = fmaf(pi, pi, wn);
        un = fmaf(u, u, un);
    }
    wn = dsb::block_sum(wn, red);
    un = dsb::block_sum(un, red);
    if (threadIdx.x == 0) {
        partials[2 * blockIdx.x] = wn;
        partials[2 * blockIdx.x + 1] = un;
    }
}

__global__ void __launch_bounds__(512) lamb_phase2(float* __restrict__ p, const float* __restrict__ upd, void* lp,
                                                   int lpdt, int64_t n, const float* __restrict__ partials,
                                                   int nblocks1, float* __restrict__ coeff_out, float lr,
                                                   float max_coeff, float min_coeff)
{
    __shared__ float red[33];
    __shared__ float s_coeff;
    float wn = 0.f, un = 0.f;
    for (int i = threadIdx.x; i < nblocks1; i += blockDim.x) {
        wn += partials[2 * i];
        un += partials[2 * i + 1];
    }
    wn = dsb::block_sum(wn, red);
    un = dsb::block_sum(un, red);
    if (threadIdx.x == 0) {
        float c = 1.f;
        if (wn > 0.f && un > 0.f) c = fminf(fmaxf(sqrtf(wn) / sqrtf(un), min_coeff), max_coeff);
        s_coeff = c;
        if (blockIdx.x == 0) {
            coeff_out[0] = sqrtf(wn);
            coeff_out[1] = sqrtf(un);
            coeff_out[2] = c;
        }
    }
    __syncthreads();
    const float step = lr * s_coeff;
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nth = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = tid; i < n; i += nth) {
        float pi = fmaf(-step, upd[i], p[i]);
        p[i] = pi;
        if (lp != nullptr) dsb::st_dt(lp, lpdt, i, pi);
    }
}

}  // namespace

extern "C" {

int dsb200_sumsq(const void* x, int64_t n, int dt, float* scratch, float* out, cudaStream_t stream)
{
    if (n <= 0) {
        DSB_CHECK(cudaMemsetAsync(out, 0, sizeof(float), stream));
        return 0;
    }
    // scratch: SUMSQ_MAX_BLOCKS partials + 1 ticket word. The ticket must be zero before the first use of a
    // scratch buffer (callers allocate it zero-filled); it wraps back to zero at the end of every launch.
    int64_t per_block = 512LL * 16;
    int blocks = (int)((n + per_block - 1) / per_block);
    if (blocks > DSB_NUM_SMS * 4) blocks = DSB_NUM_SMS * 4;
    if (blocks > SUMSQ_MAX_BLOCKS) blocks = SUMSQ_MAX_BLOCKS;
    if (blocks < 1) blocks = 1;
    if (dt == DSB_DT_F32)
        sumsq_kernel<float><<<blocks, 512, 0, stream>>>((const float*)x, n, scratch, out);
    else if (dt == DSB_DT_BF16)
        sumsq_kernel<__nv_bfloat16><<<blocks, 512, 0, stream>>>((const __nv_bfloat16*)x, n, scratch, out);
    else
        sumsq_kernel<__half><<<blocks, 512, 0, stream>>>((const __half*)x, n, scratch, out);
    DSB_CHECK_LAUNCH("sumsq");
    return 0;
}

int dsb200_multi_tensor_optim(int kind, const void* table, int n_chunks, const float* hyper_host, int n_groups,
                              float grad_scale, cudaStream_t stream)
{
    if (n_chunks <= 0) return 0;
    if (n_groups > MAX_GROUPS) {
        dsb200_set_error("multi_tensor_optim: more than 64 param groups");
        return -1;
    }
    Hyper h;
    for (int g = 0; g < n_groups; g++)
        for (int k = 0; k < 8; k++) h.h[g][k] = hyper_host[g * 8 + k];
    const Chunk* t = reinterpret_cast<const Chunk*>(table);
    switch (kind) {
        case K_ADAM: multi_tensor_optim_kernel<K_ADAM><<<n_chunks, 512, 0, stream>>>(t, h, grad_scale); break;
        case K_LION: multi_tensor_optim_kernel<K_LION><<<n_chunks, 512, 0, stream>>>(t, h, grad_scale); break;
        case K_ADAGRAD: multi_tensor_optim_kernel<K_ADAGRAD><<<n_chunks, 512, 0, stream>>>(t, h, grad_scale); break;
        case K_SGD: multi_tensor_optim_kernel<K_SGD><<<n_chunks, 512, 0, stream>>>(t, h, grad_scale); break;
        default: dsb200_set_error("multi_tensor_optim: unknown kind"); return -1;
    }
    DSB_CHECK_LAUNCH("multi_tensor_optim");
    return 0;
}

int dsb200_lamb_step(float* p, float* g, float* m, float* v, void* lp, int64_t n, int lpdt, float* partials,
                     float* coeff_out, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2,
                     int eps_mode, float max_coeff, float min_coeff, cudaStream_t stream)
{
    if (n <= 0) return 0;
    int blocks = (int)((n + 512 * 8 - 1) / (512 * 8));
    if (blocks > DSB_NUM_SMS * 4) blocks = DSB_NUM_SMS * 4;
    if (blocks > 2048) blocks = 2048;
    lamb_phase1<<<blocks, 512, 0, stream>>>(p, g, m, v, n, partials, b1, b2, eps, wd, bc1, bc2, eps_mode);
    DSB_CHECK_LAUNCH("lamb_phase1");
    lamb_phase2<<<blocks, 512, 0, stream>>>(p, g, lp, lpdt, n, partials, blocks, coeff_out, lr, max_coeff, min_coeff);
    DSB_CHECK_LAUNCH("lamb_phase2");
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// dsb200_sumsq_noalloc: the sum-of-squares reduction for callers that cannot hand us scratch memory.
//
// Who needs this: a host program that has given all of the device memory to its own pool (a serving process, a
// framework allocator that reserved the whole HBM) cannot call cudaMalloc for our 4 KB of partial sums without
// failing or forcing that pool to give memory back; a latency-critical loop must not call cudaMalloc/cudaFree
// because both synchronise the whole device; and a plain C caller may simply have no allocator to ask. For them
// the partial sums live in the library image itself: a file-scope __device__ array that every GPU which loads the
// library carries until the process exits (SUMSQ_STATIC_BYTES below, about 4.3 KB per GPU).
//
// Rules that follow from sharing one array:
//   * The name `g_sumsq_static_scratch` is only a host-side handle in host code. The device address is obtained
//     with cudaGetSymbolAddress, once per device, with that device current; this happens inside the FIRST call of
//     dsb200_sumsq_noalloc on each device. Make that first call before starting any stream capture.
//   * All calls on one GPU use the same array and it keeps whatever the previous call left. The ticket word has
//     to start at zero, so every call clears it (cudaMemsetAsync on the caller's stream) before the kernel.
//   * Two calls that overlap on the same GPU (different streams, different host threads) race on the array and
//     produce garbage. Serialise them: one stream, or an event between them.
//   * Inputs larger than DSB200_SUMSQ_NOALLOC_MAX_ELEMS are refused: beyond that the per-thread fp32
//     accumulation chains get long enough that the result no longer meets the library's accuracy bound.
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ---------------------------------------------------------------------------------------------------------
#include <mutex>

#ifndef DSB200_SUMSQ_NOALLOC_MAX_ELEMS
#define DSB200_SUMSQ_NOALLOC_MAX_ELEMS (1LL << 34)  // keep in sync with include/dsb200.h
#endif

namespace {
constexpr int SUMSQ_STATIC_FLOATS = SUMSQ_MAX_BLOCKS + 64;  // partial sums + ticket word (+ padding)
constexpr int SUMSQ_STATIC_BYTES = SUMSQ_STATIC_FLOATS * 4;
constexpr int SUMSQ_STATIC_MAX_DEVICES = 64;
static_assert(SUMSQ_STATIC_BYTES <= 8192, "the per-GPU static scratch is meant to stay a few KB");
}  // namespace

static __device__ __align__(256) float g_sumsq_static_scratch[SUMSQ_STATIC_FLOATS];

namespace {
float* g_sumsq_static_ptr[SUMSQ_STATIC_MAX_DEVICES] = {nullptr};
std::mutex g_sumsq_static_mu;

// Device address of the static scratch on the current device (looked up once per device).
int sumsq_static_address(float** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        dsb200_set_error(cudaGetErrorString(e));
        return (int)e;
    }
    if (dev < 0 || dev >= SUMSQ_STATIC_MAX_DEVICES) {
        dsb200_set_error("sumsq_noalloc: device ordinal out of range (max 64 devices)");
        return -1;
    }
    std::lock_guard<std::mutex> lk(g_sumsq_static_mu);
    if (g_sumsq_static_ptr[dev] == nullptr) {
        float* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_sumsq_static_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            dsb200_set_error(e != cudaSuccess ? cudaGetErrorString(e) : "sumsq_noalloc: null symbol address");
            return e != cudaSuccess ? (int)e : -1;
        }
        g_sumsq_static_ptr[dev] = ptr;
    }
    *out = g_sumsq_static_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int dsb200_sumsq_noalloc(const void* x, int64_t n, int dt, float* out, cudaStream_t stream)
{
    if (n > DSB200_SUMSQ_NOALLOC_MAX_ELEMS) {
        dsb200_set_error("sumsq_noalloc: n exceeds DSB200_SUMSQ_NOALLOC_MAX_ELEMS (2^34 elements); split the input");
        return -2;
    }
    if (dt != DSB_DT_F32 && dt != DSB_DT_F16 && dt != DSB_DT_BF16) {
        dsb200_set_error("sumsq_noalloc: dtype must be 0 (fp32), 1 (fp16) or 2 (bf16)");
        return -3;
    }
    if (n <= 0) {
        DSB_CHECK(cudaMemsetAsync(out, 0, sizeof(float), stream));
        return 0;
    }
    float* scratch = nullptr;
    int rc = sumsq_static_address(&scratch);
    if (rc != 0) return rc;
    // the array still holds the previous call's data: the ticket must be zero when the kernel starts
    DSB_CHECK(cudaMemsetAsync(scratch + SUMSQ_MAX_BLOCKS, 0, sizeof(unsigned int), stream));
    int64_t per_block = 512LL * 16;
    int blocks = (int)((n + per_block - 1) / per_block);
    if (blocks > DSB_NUM_SMS * 4) blocks = DSB_NUM_SMS * 4;
    if (blocks > SUMSQ_MAX_BLOCKS) blocks = SUMSQ_MAX_BLOCKS;
    if (blocks < 1) blocks = 1;
    if (dt == DSB_DT_F32)
        sumsq_kernel<float><<<blocks, 512, 0, stream>>>((const float*)x, n, scratch, out);
    else if (dt == DSB_DT_BF16)
        sumsq_kernel<__nv_bfloat16><<<blocks, 512, 0, stream>>>((const __nv_bfloat16*)x, n, scratch, out);
    else
        sumsq_kernel<__half><<<blocks, 512, 0, stream>>>((const __half*)x, n, scratch, out);
    DSB_CHECK_LAUNCH("sumsq_noalloc");
    return 0;
}
