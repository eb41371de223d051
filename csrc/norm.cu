// RMSNorm / LayerNorm forward + backward for sm_100a.
//
// Reference kernels: csrc/transformer/inference/csrc/rms_norm.cu:22,93 (fwd only, fused residual),
// csrc/transformer/normalize_kernels.cu (LayerNorm fwd + LayerNormBackward1/2). The reference has no RMSNorm
// backward at all (training falls back to eager PyTorch: 6+ elementwise passes).
// Design: one CTA per row with the whole row register-resident (single HBM read), 16-byte accesses, fp32 math;
// optional fused residual add (writes the updated residual stream). Backward: a persistent grid
// (148 SMs x 4 CTAs) loops over rows, accumulates dgamma/dbeta partials in registers and writes one partial row
// per CTA into scratch; a second tiny kernel reduces the partials (deterministic, no atomics).
#include "common.cuh"

namespace {

constexpr int MAX_VPT_LIMIT = 8;  // 16-byte vectors per thread -> up to 8*8*1024 = 65536 columns

constexpr int max_threads_for(int vpt) { return vpt == 1 ? 128 : (vpt == 8 ? 1024 : 512); }

template <typename T, bool LAYERNORM, int MAX_VPT>
__global__ void __launch_bounds__(max_threads_for(MAX_VPT)) norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ residual,
                                                        const T* __restrict__ gamma, const T* __restrict__ beta,
                                                        T* __restrict__ y, T* __restrict__ res_out,
                                                        float* __restrict__ stats, int rows, int cols, float eps)
{
    __shared__ float red[33];
    const int row = blockIdx.x;
    const int nvec = cols >> 3;
    const T* xr = x + (int64_t)row * cols;
    dsb::f8 cache[MAX_VPT];
    float sum = 0.f, sumsq = 0.f;
#pragma unroll
    for (int k = 0; k < MAX_VPT; k++) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            dsb::f8 a = dsb::load8(xr + v * 8);
            if (residual != nullptr) {
                dsb::f8 r = dsb::load8(residual + (int64_t)row * cols + v * 8);
#pragma unroll
                for (int i = 0; i < 8; i++) a.v[i] += r.v[i];
                if (res_out != nullptr) {
                    dsb::store8(res_out + (int64_t)row * cols + v * 8, a);
                    // keep numerics identical to "norm(bf16(x+res))": round through the storage type
#pragma unroll
                    for (int i = 0; i < 8; i++) a.v[i] = dsb::to_f<T>(dsb::from_f<T>(a.v[i]));
                }
            }
            cache[k] = a;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                sum += a.v[i];
                sumsq = fmaf(a.v[i], a.v[i], sumsq);
            }
        }
    }
    float mean = 0.f, rstd;
    if (LAYERNORM) {
        mean = dsb::block_sum(sum, red) / cols;
        // two-pass variance from the register cache (numerically safer than E[x^2]-mean^2)
        float var = 0.f;
#pragma unroll
        for (int k = 0; k < MAX_VPT; k++) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    float d = cache[k].v[i] - mean;
                    var = fmaf(d, d, var);
                }
            }
        }
        var = dsb::block_sum(var, red) / cols;
        rstd = rsqrtf(var + eps);
    } else {
        float ms = dsb::block_sum(sumsq, red) / cols;
        rstd = rsqrtf(ms + eps);
    }
    if (threadIdx.x == 0 && stats != nullptr) {
        if (LAYERNORM) {
            stats[2 * row] = mean;
            stats[2 * row + 1] = rstd;
        } else {
            stats[row] = rstd;
        }
    }
#pragma unroll
    for (int k = 0; k < MAX_VPT; k++) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            dsb::f8 g = dsb::load8(gamma + v * 8);
            dsb::f8 o;
            if (LAYERNORM && beta != nullptr) {
                dsb::f8 b = dsb::load8(beta + v * 8);
#pragma unroll
                for (int i = 0; i < 8; i++) o.v[i] = fmaf((cache[k].v[i] - mean) * rstd, g.v[i], b.v[i]);
            } else {
#pragma unroll
                for (int i = 0; i < 8; i++) o.v[i] = (cache[k].v[i] - mean) * rstd * g.v[i];
            }
            dsb::store8(y + (int64_t)row * cols + v * 8, o);
        }
    }
}

// Backward. dx = rstd * (dy*g - xhat * mean(dy*g*xhat) [- mean(dy*g) for LN]); dgamma += dy*xhat; dbeta += dy.
// If dres != nullptr the incoming residual-stream gradient is added to dx (fused residual backward).
template <typename T, bool LAYERNORM, int MAX_VPT>
__global__ void __launch_bounds__(max_threads_for(MAX_VPT)) norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                        const T* __restrict__ gamma, const float* __restrict__ stats,
                                                        const T* __restrict__ dres, T* __restrict__ dx,
                                                        float* __restrict__ partial_dg, float* __restrict__ partial_db,
                                                        int rows, int cols)
{
    __shared__ float red[33];
    const int nvec = cols >> 3;
    dsb::f8 gam[MAX_VPT], dg[MAX_VPT], db[MAX_VPT];
#pragma unroll
    for (int k = 0; k < MAX_VPT; k++) {
        const int v = threadIdx.x + k * blockDim.x;
#pragma unroll
        for (int i = 0; i < 8; i++) dg[k].v[i] = db[k].v[i] = 0.f;
        if (v < nvec) gam[k] = dsb::load8(gamma + v * 8);
    }
    // software pipeline over rows: the 16-byte loads of the NEXT row are issued before the block reduction of the
    // current one, so the DRAM latency of row r+1 hides behind the reduction + stores of row r
    uint4 rx[MAX_VPT], rdy[MAX_VPT], rres[MAX_VPT];
    auto fetch = [&](int row) {
#pragma unroll
        for (int k = 0; k < MAX_VPT; k++) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
                rx[k] = dsb::ldg_stream(reinterpret_cast<const uint4*>(x + (int64_t)row * cols) + v);
                rdy[k] = dsb::ldg_stream(reinterpret_cast<const uint4*>(dy + (int64_t)row * cols) + v);
                if (dres != nullptr) rres[k] = dsb::ldg_stream(reinterpret_cast<const uint4*>(dres + (int64_t)row * cols) + v);
            }
        }
    };
    constexpr bool PIPE = MAX_VPT <= 2;  // wider rows would spill the extra row of registers
    if (PIPE && (int)blockIdx.x < rows) fetch(blockIdx.x);
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        if (!PIPE) fetch(row);
        const float mean = LAYERNORM ? stats[2 * row] : 0.f;
        const float rstd = LAYERNORM ? stats[2 * row + 1] : stats[row];
        dsb::f8 xh[MAX_VPT], dyg[MAX_VPT];
        uint4 cres[MAX_VPT];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int k = 0; k < MAX_VPT; k++) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
                dsb::f8 a = dsb::load8(reinterpret_cast<const T*>(&rx[k]));
                dsb::f8 d = dsb::load8(reinterpret_cast<const T*>(&rdy[k]));
                if (PIPE) cres[k] = rres[k];
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    float h = (a.v[i] - mean) * rstd;
                    float t = d.v[i] * gam[k].v[i];
                    xh[k].v[i] = h;
                    dyg[k].v[i] = t;
                    dg[k].v[i] = fmaf(d.v[i], h, dg[k].v[i]);
                    db[k].v[i] += d.v[i];
                    s1 = fmaf(t, h, s1);
                    s2 += t;
                }
            }
        }
        if (PIPE && row + (int)gridDim.x < rows) fetch(row + gridDim.x);
        s1 = dsb::block_sum(s1, red) / cols;
        s2 = LAYERNORM ? dsb::block_sum(s2, red) / cols : 0.f;
#pragma unroll
        for (int k = 0; k < MAX_VPT; k++) {
            const int v = threadIdx.x + k * blockDim.x;
            if (v < nvec) {
                dsb::f8 o;
#pragma unroll
                for (int i = 0; i < 8; i++) o.v[i] = rstd * (dyg[k].v[i] - xh[k].v[i] * s1 - s2);
                if (dres != nullptr) {
                    dsb::f8 r = dsb::load8(reinterpret_cast<const T*>(PIPE ? &cres[k] : &rres[k]));
#pragma unroll
                    for (int i = 0; i < 8; i++) o.v[i] += r.v[i];
                }
                dsb::store8(dx + (int64_t)row * cols + v * 8, o);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < MAX_VPT; k++) {
        const int v = threadIdx.x + k * blockDim.x;
        if (v < nvec) {
            float* pg = partial_dg + (int64_t)blockIdx.x * cols + v * 8;
            *reinterpret_cast<float4*>(pg) = make_float4(dg[k].v[0], dg[k].v[1], dg[k].v[2], dg[k].v[3]);
            *reinterpret_cast<float4*>(pg + 4) = make_float4(dg[k].v[4], dg[k].v[5], dg[k].v[6], dg[k].v[7]);
            if (LAYERNORM && partial_db != nullptr) {
                float* pb = partial_db + (int64_t)blockIdx.x * cols + v * 8;
                *reinterpret_cast<float4*>(pb) = make_float4(db[k].v[0], db[k].v[1], db[k].v[2], db[k].v[3]);
                *reinterpret_cast<float4*>(pb + 4) = make_float4(db[k].v[4], db[k].v[5], db[k].v[6], db[k].v[7]);
            }
        }
    }
}

// out[c] (T or float, accumulate optional) = sum_b partial[b][c]
template <typename T>
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float* __restrict__ partial, int nparts, int cols,
                                                              T* __restrict__ out, int accumulate)
{
    // 32 columns x 8 row groups per block: every warp reads one 128-byte line per partial row
    __shared__ float sm[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float s = 0.f;
    if (c < cols)
        for (int b = ty; b < nparts; b += 8) s += partial[(int64_t)b * cols + c];
    sm[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < cols) {
#pragma unroll
        for (int j = 1; j < 8; j++) s += sm[j][tx];
        if (accumulate) s += dsb::to_f<T>(out[c]);
        out[c] = dsb::from_f<T>(s);
    }
}

int pick_vpt(int cols)
{
    int nvec = cols / 8;
    if (nvec <= 128) return 1;   // <= 128 threads
    if (nvec <= 1024) return 2;  // <= 512 threads
    if (nvec <= 2048) return 4;  // <= 512 threads
    return 8;                    // <= 1024 threads
}
int pick_threads(int cols, int vpt)
{
    int nvec = cols / 8;
    int t = (nvec + vpt - 1) / vpt;
    if (t < 32) t = 32;
    t = (t + 31) / 32 * 32;
    return t > 1024 ? 1024 : t;
}

#define DSB_VPT_SWITCH(vpt, ...)                      \
    switch (vpt) {                                    \
        case 1: { constexpr int VPT = 1; __VA_ARGS__; } break; \
        case 2: { constexpr int VPT = 2; __VA_ARGS__; } break; \
        case 4: { constexpr int VPT = 4; __VA_ARGS__; } break; \
        default: { constexpr int VPT = 8; __VA_ARGS__; } break; \
    }

template <typename T>
int launch_fwd(bool ln, const void* x, const void* res, const void* gamma, const void* beta, void* y, void* res_out,
               float* stats, int rows, int cols, float eps, cudaStream_t s)
{
    const int vpt = pick_vpt(cols), th = pick_threads(cols, vpt);
    DSB_VPT_SWITCH(vpt, {
        if (ln)
            norm_fwd_kernel<T, true, VPT><<<rows, th, 0, s>>>((const T*)x, (const T*)res, (const T*)gamma,
                                                              (const T*)beta, (T*)y, (T*)res_out, stats, rows, cols, eps);
        else
            norm_fwd_kernel<T, false, VPT><<<rows, th, 0, s>>>((const T*)x, (const T*)res, (const T*)gamma,
                                                               (const T*)beta, (T*)y, (T*)res_out, stats, rows, cols, eps);
    });
    DSB_CHECK_LAUNCH("norm_fwd");
    return 0;
}

template <typename T>
int launch_bwd(bool ln, const void* dy, const void* x, const void* gamma, const float* stats, const void* dres, void* dx,
               void* dgamma, void* dbeta, float* scratch, int rows, int cols, int accumulate, cudaStream_t s)
{
    const int vpt = pick_vpt(cols), th = pick_threads(cols, vpt);
    int grid = DSB_NUM_SMS * (th <= 256 ? 4 : (th <= 512 ? 2 : 1));
    if (grid > rows) grid = rows;
    float* pdg = scratch;
    float* pdb = scratch + (int64_t)grid * cols;
    DSB_VPT_SWITCH(vpt, {
        if (ln)
            norm_bwd_kernel<T, true, VPT><<<grid, th, 0, s>>>((const T*)dy, (const T*)x, (const T*)gamma, stats,
                                                              (const T*)dres, (T*)dx, pdg, pdb, rows, cols);
        else
            norm_bwd_kernel<T, false, VPT><<<grid, th, 0, s>>>((const T*)dy, (const T*)x, (const T*)gamma, stats,
                                                               (const T*)dres, (T*)dx, pdg, nullptr, rows, cols);
    });
    DSB_CHECK_LAUNCH("norm_bwd");
    reduce_partials_kernel<T><<<(cols + 31) / 32, 256, 0, s>>>(pdg, grid, cols, (T*)dgamma, accumulate);
    DSB_CHECK_LAUNCH("norm_bwd_reduce_g");
    if (ln && dbeta != nullptr) {
        reduce_partials_kernel<T><<<(cols + 31) / 32, 256, 0, s>>>(pdb, grid, cols, (T*)dbeta, accumulate);
        DSB_CHECK_LAUNCH("norm_bwd_reduce_b");
    }
    return 0;
}

bool shape_ok(int cols)
{
    if (cols % 8 != 0 || cols / 8 > MAX_VPT_LIMIT * 1024) {
        dsb200_set_error("norm: cols must be a multiple of 8 and <= 65536");
        return false;
    }
    return true;
}

}  // namespace

extern "C" {

// stats: fp32 [rows] (rstd). residual/res_out may be null. dt: 1=fp16 2=bf16.
int dsb200_rmsnorm_fwd(const void* x, const void* residual, const void* gamma, void* y, void* res_out, float* stats,
                         int rows, int cols, float eps, int dt, cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (!shape_ok(cols)) return -1;
    if (dt == DSB_DT_BF16)
        return launch_fwd<__nv_bfloat16>(false, x, residual, gamma, nullptr, y, res_out, stats, rows, cols, eps, stream);
    if (dt == DSB_DT_F16)
        return launch_fwd<__half>(false, x, residual, gamma, nullptr, y, res_out, stats, rows, cols, eps, stream);
    dsb200_set_error("rmsnorm: dtype must be fp16/bf16");
    return -1;
}

// scratch: fp32 [grid*cols] with grid <= 148*4.
int dsb200_rmsnorm_bwd(const void* dy, const void* x, const void* gamma, const float* stats, const void* dres,
                         void* dx, void* dgamma, float* scratch, int rows, int cols, int accumulate, int dt,
                         cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (!shape_ok(cols)) return -1;
    if (dt == DSB_DT_BF16)
        return launch_bwd<__nv_bfloat16>(false, dy, x, gamma, stats, dres, dx, dgamma, nullptr, scratch, rows, cols,
                                         accumulate, stream);
    if (dt == DSB_DT_F16)
        return launch_bwd<__half>(false, dy, x, gamma, stats, dres, dx, dgamma, nullptr, scratch, rows, cols,
                                  accumulate, stream);
    dsb200_set_error("rmsnorm: dtype must be fp16/bf16");
    return -1;
}

// stats: fp32 [rows*2] (mean, rstd)
int dsb200_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                           void* res_out, float* stats, int rows, int cols, float eps, int dt, cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (!shape_ok(cols)) return -1;
    if (dt == DSB_DT_BF16)
        return launch_fwd<__nv_bfloat16>(true, x, residual, gamma, beta, y, res_out, stats, rows, cols, eps, stream);
    if (dt == DSB_DT_F16)
        return launch_fwd<__half>(true, x, residual, gamma, beta, y, res_out, stats, rows, cols, eps, stream);
    dsb200_set_error("layernorm: dtype must be fp16/bf16");
    return -1;
}

// scratch: fp32 [2*grid*cols]
int dsb200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* stats, const void* dres,
                           void* dx, void* dgamma, void* dbeta, float* scratch, int rows, int cols, int accumulate,
                           int dt, cudaStream_t stream)
{
    if (rows <= 0) return 0;
    if (!shape_ok(cols)) return -1;
    if (dt == DSB_DT_BF16)
        return launch_bwd<__nv_bfloat16>(true, dy, x, gamma, stats, dres, dx, dgamma, dbeta, scratch, rows, cols,
                                         accumulate, stream);
    if (dt == DSB_DT_F16)
        return launch_bwd<__half>(true, dy, x, gamma, stats, dres, dx, dgamma, dbeta, scratch, rows, cols, accumulate,
                                  stream);
    dsb200_set_error("layernorm: dtype must be fp16/bf16");
    return -1;
}

}  // extern "C"
