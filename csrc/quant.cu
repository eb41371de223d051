// Quantization kernels for ZeRO++ (qwZ / qgZ) and the FP quantizer.
//
// Reference kernels: csrc/quantization/quantize.cu:23 (cached_quantization: per-group symmetric/asymmetric int8/int4),
// dequantize.cu:12, swizzled_quantize.cu:22-83 (qgZ stage 1: quantize + [node,dev]->[dev,node] re-order),
// quant_reduce.cu:21-133 (qgZ stage 2: dequantize N chunks, sum, re-quantize), csrc/fp_quantizer/fp_quantize_impl.cu
// (group-wise minifloat quantization by bit-twiddling, one warp per group).
// B200 design:
//   * the native wire format is block-scaled FP8 (e4m3, one fp32 scale per 32..512-element block) produced with the
//     hardware converter cvt.rn.satfinite.e4m3x2.f32 - no bit twiddling; e5m2 and e2m1 (fp4) use the same path;
//   * int8/int4 symmetric group quantization is kept for bit-parity with reference checkpoints/algorithms;
//   * every kernel is one pass over HBM with 16-byte accesses, one warp per block so the abs-max is a shuffle
//     reduction and the payload + scale are written in the same pass;
//   * dequant-reduce (qgZ) accumulates in fp32 and can either re-quantize (multi-node semantics) or emit bf16/fp32.
#include "common.cuh"
#include <cuda_fp8.h>

namespace {

__device__ __forceinline__ uint16_t cvt_e4m3x2(float lo, float hi)
{
    uint16_t r;
    asm("cvt.rn.satfinite.e4m3x2.f32 %0, %2, %1;" : "=h"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ uint16_t cvt_e5m2x2(float lo, float hi)
{
    uint16_t r;
    asm("cvt.rn.satfinite.e5m2x2.f32 %0, %2, %1;" : "=h"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float2 cvt_e4m3x2_to_f2(uint16_t v)
{
    uint32_t h2;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(v));
    return __half22float2(*reinterpret_cast<__half2*>(&h2));
}
__device__ __forceinline__ float2 cvt_e5m2x2_to_f2(uint16_t v)
{
    uint32_t h2;
    asm("cvt.rn.f16x2.e5m2x2 %0, %1;" : "=r"(h2) : "h"(v));
    return __half22float2(*reinterpret_cast<__half2*>(&h2));
}

// ---------------------------------------------------------------------------------------------------------
// FP8 block quantization. One warp per block of `block` elements (multiple of 256 = 32 lanes x 8).
// q: uint8 [n], scales: fp32 [n / block]; x = q * scale. fmt: 0 = e4m3 (max 448), 1 = e5m2 (max 57344).
// ---------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) fp8_block_quant_kernel(const T* __restrict__ x, uint8_t* __restrict__ q,
                                                              float* __restrict__ scales, int64_t nblocks, int block,
                                                              int fmt)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const float fmax = fmt == 0 ? 448.f : 57344.f;
    const int iters = block / 256;
    for (int64_t b = warp; b < nblocks; b += nwarps) {
        const T* xb = x + b * block;
        float amax = 0.f;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xb + (it * 32 + lane) * 8);
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v.v[i]));
        }
        amax = dsb::warp_max(amax);
        const float scale = amax > 0.f ? amax / fmax : 1.f;
        const float inv = 1.f / scale;
        if (lane == 0) scales[b] = scale;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xb + (it * 32 + lane) * 8);  // second read hits L1/L2
            uint16_t p[4];
#pragma unroll
            for (int i = 0; i < 4; i++)
                p[i] = fmt == 0 ? cvt_e4m3x2(v.v[2 * i] * inv, v.v[2 * i + 1] * inv)
                                : cvt_e5m2x2(v.v[2 * i] * inv, v.v[2 * i + 1] * inv);
            uint2 out;
            out.x = (uint32_t)p[0] | ((uint32_t)p[1] << 16);
            out.y = (uint32_t)p[2] | ((uint32_t)p[3] << 16);
            *reinterpret_cast<uint2*>(q + b * block + (it * 32 + lane) * 8) = out;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) fp8_block_dequant_kernel(const uint8_t* __restrict__ q,
                                                                const float* __restrict__ scales, T* __restrict__ y,
                                                                int64_t n, int block, int fmt)
{
    const int64_t nvec = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        const uint2 raw = *reinterpret_cast<const uint2*>(q + i * 8);
        const float s = scales[(i * 8) / block];
        const uint16_t p[4] = {(uint16_t)(raw.x & 0xffff), (uint16_t)(raw.x >> 16), (uint16_t)(raw.y & 0xffff),
                               (uint16_t)(raw.y >> 16)};
        dsb::f8 o;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float2 f = fmt == 0 ? cvt_e4m3x2_to_f2(p[k]) : cvt_e5m2x2_to_f2(p[k]);
            o.v[2 * k] = f.x * s;
            o.v[2 * k + 1] = f.y * s;
        }
        dsb::store8(y + i * 8, o);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Symmetric integer group quantization (int8, or int4 packed two per byte, low nibble first), one warp per
// group, group size multiple of 256. scale = amax / qmax ; q = round(x / scale).
// Optional `swizzle`: output group g is written at position perm(g) so that a following all-to-all over
// [nodes, devices] sees contiguous per-destination chunks (qgZ stage 1).
// ---------------------------------------------------------------------------------------------------------
template <typename T, int BITS>
__global__ void __launch_bounds__(256) int_group_quant_kernel(const T* __restrict__ x, int8_t* __restrict__ q,
                                                              float* __restrict__ scales, int64_t ngroups, int group,
                                                              int swz_nodes, int swz_devs)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr float qmax = BITS == 8 ? 127.f : 7.f;
    const int iters = group / 256;
    const int64_t parts = (int64_t)swz_nodes * swz_devs;
    const int64_t groups_per_part = parts > 0 ? ngroups / parts : 0;
    for (int64_t g = warp; g < ngroups; g += nwarps) {
        const T* xg = x + g * group;
        float amax = 0.f;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xg + (it * 32 + lane) * 8);
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v.v[i]));
        }
        amax = dsb::warp_max(amax);
        const float scale = amax > 0.f ? amax / qmax : 1.f;
        const float inv = 1.f / scale;
        int64_t og = g;
        if (parts > 0) {  // [node, dev, j] -> [dev, node, j]
            const int64_t part = g / groups_per_part, j = g % groups_per_part;
            const int64_t node = part / swz_devs, dev = part % swz_devs;
            og = (dev * swz_nodes + node) * groups_per_part + j;
        }
        if (lane == 0) scales[og] = scale;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xg + (it * 32 + lane) * 8);
            int qi[8];
#pragma unroll
            for (int i = 0; i < 8; i++) qi[i] = max(-(int)qmax, min((int)qmax, __float2int_rn(v.v[i] * inv)));
            if (BITS == 8) {
                uint2 out;
                out.x = (qi[0] & 0xff) | ((qi[1] & 0xff) << 8) | ((qi[2] & 0xff) << 16) | ((uint32_t)(qi[3] & 0xff) << 24);
                out.y = (qi[4] & 0xff) | ((qi[5] & 0xff) << 8) | ((qi[6] & 0xff) << 16) | ((uint32_t)(qi[7] & 0xff) << 24);
                *reinterpret_cast<uint2*>(q + og * group + (it * 32 + lane) * 8) = out;
            } else {
                uint32_t out = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) out |= (uint32_t)(qi[i] & 0xf) << (4 * i);
                *reinterpret_cast<uint32_t*>(q + (og * group >> 1) + (it * 32 + lane) * 4) = out;
            }
        }
    }
}

template <typename T, int BITS>
__global__ void __launch_bounds__(256) int_group_dequant_kernel(const int8_t* __restrict__ q,
                                                                const float* __restrict__ scales, T* __restrict__ y,
                                                                int64_t n, int group)
{
    const int64_t nvec = n >> 3;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        const float s = scales[(i * 8) / group];
        dsb::f8 o;
        if (BITS == 8) {
            const uint2 raw = *reinterpret_cast<const uint2*>(q + i * 8);
            const int8_t* b = reinterpret_cast<const int8_t*>(&raw);
#pragma unroll
            for (int k = 0; k < 8; k++) o.v[k] = (float)b[k] * s;
        } else {
            const uint32_t raw = *reinterpret_cast<const uint32_t*>(q + i * 4);
#pragma unroll
            for (int k = 0; k < 8; k++) {
                int v = (raw >> (4 * k)) & 0xf;
                v = (v ^ 8) - 8;  // sign-extend the nibble
                o.v[k] = (float)v * s;
            }
        }
        dsb::store8(y + i * 8, o);
    }
}

// ---------------------------------------------------------------------------------------------------------
// LoCo (error-feedback) variant of the group quantizer (reference swizzled_quantize.cu:200-316
// loco_swizzled_quant_kernel): the value that is quantized is v = x + beta * err, and the error buffer is replaced by
// what the quantization lost, err = v - dequant(q). `err` is fp32, laid out like x (unswizzled).
// ---------------------------------------------------------------------------------------------------------
template <typename T, int BITS>
__global__ void __launch_bounds__(256) loco_group_quant_kernel(const T* __restrict__ x, float* __restrict__ err,
                                                               int8_t* __restrict__ q, float* __restrict__ scales,
                                                               int64_t ngroups, int group, float beta, int swz_nodes,
                                                               int swz_devs)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    constexpr float qmax = BITS == 8 ? 127.f : 7.f;
    const int iters = group / 256;
    const int64_t parts = (int64_t)swz_nodes * swz_devs;
    const int64_t groups_per_part = parts > 0 ? ngroups / parts : 0;
    for (int64_t g = warp; g < ngroups; g += nwarps) {
        const T* xg = x + g * group;
        float* eg = err + g * group;
        float amax = 0.f;
        for (int it = 0; it < iters; it++) {
            const int o = (it * 32 + lane) * 8;
            dsb::f8 v = dsb::load8(xg + o);
            const float4 e0 = *reinterpret_cast<const float4*>(eg + o), e1 = *reinterpret_cast<const float4*>(eg + o + 4);
            const float ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v.v[i] + beta * ev[i]));
        }
        amax = dsb::warp_max(amax);
        const float scale = amax > 0.f ? amax / qmax : 1.f;
        const float inv = 1.f / scale;
        int64_t og = g;
        if (parts > 0) {
            const int64_t part = g / groups_per_part, j = g % groups_per_part;
            const int64_t node = part / swz_devs, dev = part % swz_devs;
            og = (dev * swz_nodes + node) * groups_per_part + j;
        }
        if (lane == 0) scales[og] = scale;
        for (int it = 0; it < iters; it++) {
            const int o = (it * 32 + lane) * 8;
            dsb::f8 v = dsb::load8(xg + o);
            const float4 e0 = *reinterpret_cast<const float4*>(eg + o), e1 = *reinterpret_cast<const float4*>(eg + o + 4);
            const float ev[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
            int qi[8];
            float ne[8];
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float t = v.v[i] + beta * ev[i];
                qi[i] = max(-(int)qmax, min((int)qmax, __float2int_rn(t * inv)));
                ne[i] = t - (float)qi[i] * scale;
            }
            *reinterpret_cast<float4*>(eg + o) = make_float4(ne[0], ne[1], ne[2], ne[3]);
            *reinterpret_cast<float4*>(eg + o + 4) = make_float4(ne[4], ne[5], ne[6], ne[7]);
            if (BITS == 8) {
                uint2 out;
                out.x = (qi[0] & 0xff) | ((qi[1] & 0xff) << 8) | ((qi[2] & 0xff) << 16) | ((uint32_t)(qi[3] & 0xff) << 24);
                out.y = (qi[4] & 0xff) | ((qi[5] & 0xff) << 8) | ((qi[6] & 0xff) << 16) | ((uint32_t)(qi[7] & 0xff) << 24);
                *reinterpret_cast<uint2*>(q + og * group + o) = out;
            } else {
                uint32_t out = 0;
#pragma unroll
                for (int i = 0; i < 8; i++) out |= (uint32_t)(qi[i] & 0xf) << (4 * i);
                *reinterpret_cast<uint32_t*>(q + (og * group >> 1) + (it * 32 + lane) * 4) = out;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// MoQ fake quantization in place (reference fake_quantizer.cu: fake_quantize_kernel / _asym and the stochastic-rounding
// sr_ variants): every group is snapped to its `bits`-bit grid and written back in the input dtype. One warp per group,
// group size a multiple of 256. Symmetric: q in [-2^(b-1), 2^(b-1)-1], scale = amax / 2^(b-1); asymmetric: q in
// [0, 2^b - 1] over [min, max]. Stochastic rounding adds a uniform draw in [0, 1) before the floor.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float uniform01(uint32_t seed, uint64_t idx)
{
    uint64_t z = idx + 0x9E3779B97F4A7C15ull * (uint64_t)(seed + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    return (float)(z >> 40) * (1.0f / 16777216.0f);
}

template <typename T>
__global__ void __launch_bounds__(256) fake_quant_kernel(T* __restrict__ x, int64_t ngroups, int group, int bits,
                                                         int symmetric, int stochastic, uint32_t seed)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int iters = group / 256;
    const float levels = (float)(1 << bits);
    for (int64_t g = warp; g < ngroups; g += nwarps) {
        T* xg = x + g * group;
        float mx = -INFINITY, mn = INFINITY;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xg + (it * 32 + lane) * 8);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                mx = fmaxf(mx, v.v[i]);
                mn = fminf(mn, v.v[i]);
            }
        }
        mx = dsb::warp_max(mx);
        mn = -dsb::warp_max(-mn);
        float scale, zero, lo, hi;
        if (symmetric) {
            const float amax = fmaxf(fabsf(mx), fabsf(mn));
            scale = amax > 0.f ? 2.f * amax / levels : 1.f;
            zero = 0.f;
            lo = -levels * 0.5f;
            hi = levels * 0.5f - 1.f;
        } else {
            scale = mx > mn ? (mx - mn) / (levels - 1.f) : 1.f;
            zero = mn;
            lo = 0.f;
            hi = levels - 1.f;
        }
        const float inv = 1.f / scale;
        for (int it = 0; it < iters; it++) {
            const int o = (it * 32 + lane) * 8;
            dsb::f8 v = dsb::load8(xg + o);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float t = (v.v[i] - zero) * inv;
                float qv = stochastic ? floorf(t + uniform01(seed, (uint64_t)(g * group + o + i))) : rintf(t);
                qv = fminf(hi, fmaxf(lo, qv));
                v.v[i] = qv * scale + zero;
            }
            dsb::store8(xg + o, v);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// qgZ stage 2: out = sum_c dequant(q[c]) for `nchunks` received chunks of `chunk_elems` each (int8 or int4 groups
// of size `group`), written as T (bf16/fp32-as-bf16 pairs) without re-quantization (single NVSwitch domain), or
// re-quantized to int4 groups by the caller with int_group_quant.
// ---------------------------------------------------------------------------------------------------------
template <typename T, int BITS>
__global__ void __launch_bounds__(256) dequant_reduce_kernel(const int8_t* __restrict__ q,
                                                             const float* __restrict__ scales, T* __restrict__ out,
                                                             int64_t chunk_elems, int nchunks, int group, float post)
{
    const int64_t nvec = chunk_elems >> 3;
    const int64_t groups_per_chunk = chunk_elems / group;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int c = 0; c < nchunks; c++) {
            const float s = scales[c * groups_per_chunk + (i * 8) / group];
            if (BITS == 8) {
                const uint2 raw = *reinterpret_cast<const uint2*>(q + c * chunk_elems + i * 8);
                const int8_t* b = reinterpret_cast<const int8_t*>(&raw);
#pragma unroll
                for (int k = 0; k < 8; k++) acc[k] = fmaf((float)b[k], s, acc[k]);
            } else {
                const uint32_t raw = *reinterpret_cast<const uint32_t*>(q + ((c * chunk_elems) >> 1) + i * 4);
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    int v = (raw >> (4 * k)) & 0xf;
                    v = (v ^ 8) - 8;
                    acc[k] = fmaf((float)v, s, acc[k]);
                }
            }
        }
        dsb::f8 o;
#pragma unroll
        for (int k = 0; k < 8; k++) o.v[k] = acc[k] * post;
        dsb::store8(out + i * 8, o);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Generic minifloat fake-format quantizer for the FP quantizer op: q_bits in {8 (e4m3), 6 (e3m2), 4 (e2m1), 12
// (e4m7)}: values are scaled per group to the format's max, rounded to `man_bits` mantissa bits (RNE or
// stochastic), and stored as scaled e4m3 bytes (8-bit), or as bf16-carried codes for 6/12 bit (the packing of 6/12
// bit payloads is done by the Python wrapper with integer ops). Output here: the dequantized-rounded value / scale
// in T (i.e. the exact representable number), plus the group scale.
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float round_to_minifloat(float x, int exp_bits, int man_bits, float maxv, uint32_t rnd,
                                                    bool stochastic)
{
    if (x == 0.f) return 0.f;
    const float ax = fminf(fabsf(x), maxv);
    int e;
    frexpf(ax, &e);  // ax = m * 2^e, m in [0.5,1)
    const int bias = (1 << (exp_bits - 1)) - 1;
    int ee = e - 1;  // exponent of the leading 1
    if (ee < 1 - bias) ee = 1 - bias;  // subnormal range shares the smallest exponent
    const float ulp = ldexpf(1.f, ee - man_bits);
    float qv;
    if (stochastic) {
        const float r = (float)(rnd & 0xffffff) / 16777216.f;
        qv = floorf(ax / ulp + r) * ulp;
    } else {
        qv = rintf(ax / ulp) * ulp;
    }
    qv = fminf(qv, maxv);
    return copysignf(qv, x);
}

template <typename T>
__global__ void __launch_bounds__(256) minifloat_quant_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                              float* __restrict__ scales, int64_t ngroups, int group,
                                                              int exp_bits, int man_bits, float fmt_max,
                                                              int stochastic, uint32_t seed)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int iters = group / 256;
    for (int64_t g = warp; g < ngroups; g += nwarps) {
        const T* xg = x + g * group;
        float amax = 0.f;
        for (int it = 0; it < iters; it++) {
            dsb::f8 v = dsb::load8(xg + (it * 32 + lane) * 8);
#pragma unroll
            for (int i = 0; i < 8; i++) amax = fmaxf(amax, fabsf(v.v[i]));
        }
        amax = dsb::warp_max(amax);
        const float scale = amax > 0.f ? amax / fmt_max : 1.f;
        const float inv = 1.f / scale;
        if (lane == 0) scales[g] = scale;
        for (int it = 0; it < iters; it++) {
            const int64_t base = (it * 32 + lane) * 8;
            dsb::f8 v = dsb::load8(xg + base), o;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                uint32_t h = (uint32_t)((g * group + base + i) * 2654435761u) ^ seed;
                h ^= h >> 15;
                h *= 2246822519u;
                h ^= h >> 13;
                o.v[i] = round_to_minifloat(v.v[i] * inv, exp_bits, man_bits, fmt_max, h, stochastic != 0);
            }
            dsb::store8(y + g * group + base, o);
        }
    }
}

int grid_warps(int64_t nwarps_needed)
{
    int64_t blocks = (nwarps_needed + 7) / 8;  // 8 warps per 256-thread block
    const int64_t cap = (int64_t)DSB_NUM_SMS * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}
int grid_elems(int64_t nvec)
{
    int64_t blocks = (nvec + 255) / 256;
    const int64_t cap = (int64_t)DSB_NUM_SMS * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

}  // namespace

extern "C" {

int dsb200_fp8_block_quant(const void* x, int dt, void* q, float* scales, int64_t n, int block, int fmt,
                           cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (block % 256 != 0 || n % block != 0) {
        dsb200_set_error("fp8_block_quant: block must be a multiple of 256 and divide n");
        return -1;
    }
    const int64_t nb = n / block;
    if (dt == DSB_DT_BF16)
        fp8_block_quant_kernel<__nv_bfloat16><<<grid_warps(nb), 256, 0, stream>>>((const __nv_bfloat16*)x, (uint8_t*)q,
                                                                                  scales, nb, block, fmt);
    else if (dt == DSB_DT_F16)
        fp8_block_quant_kernel<__half><<<grid_warps(nb), 256, 0, stream>>>((const __half*)x, (uint8_t*)q, scales, nb,
                                                                           block, fmt);
    else {
        dsb200_set_error("fp8_block_quant: input must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("fp8_block_quant");
    return 0;
}

int dsb200_fp8_block_dequant(const void* q, const float* scales, void* y, int dt, int64_t n, int block, int fmt,
                             cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (n % 8 != 0) {
        dsb200_set_error("fp8_block_dequant: n must be a multiple of 8");
        return -1;
    }
    if (dt == DSB_DT_BF16)
        fp8_block_dequant_kernel<__nv_bfloat16><<<grid_elems(n / 8), 256, 0, stream>>>((const uint8_t*)q, scales,
                                                                                       (__nv_bfloat16*)y, n, block, fmt);
    else if (dt == DSB_DT_F16)
        fp8_block_dequant_kernel<__half><<<grid_elems(n / 8), 256, 0, stream>>>((const uint8_t*)q, scales, (__half*)y, n,
                                                                                block, fmt);
    else {
        dsb200_set_error("fp8_block_dequant: output must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("fp8_block_dequant");
    return 0;
}

int dsb200_int_group_quant(const void* x, int dt, void* q, float* scales, int64_t n, int group, int bits, int swz_nodes,
                           int swz_devs, cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (group % 256 != 0 || n % group != 0 || (bits != 8 && bits != 4)) {
        dsb200_set_error("int_group_quant: group must be a multiple of 256 dividing n; bits in {4,8}");
        return -1;
    }
    const int64_t ng = n / group;
    if (swz_nodes * swz_devs > 0 && ng % ((int64_t)swz_nodes * swz_devs) != 0) {
        dsb200_set_error("int_group_quant: groups must divide evenly over nodes*devices for the swizzled layout");
        return -1;
    }
#define IQ(T, B)                                                                                                      \
    int_group_quant_kernel<T, B><<<grid_warps(ng), 256, 0, stream>>>((const T*)x, (int8_t*)q, scales, ng, group,      \
                                                                     swz_nodes, swz_devs)
    if (dt == DSB_DT_BF16) {
        if (bits == 8) IQ(__nv_bfloat16, 8); else IQ(__nv_bfloat16, 4);
    } else if (dt == DSB_DT_F16) {
        if (bits == 8) IQ(__half, 8); else IQ(__half, 4);
    } else {
        dsb200_set_error("int_group_quant: input must be fp16/bf16");
        return -1;
    }
#undef IQ
    DSB_CHECK_LAUNCH("int_group_quant");
    return 0;
}

// LoCo error-feedback quantization: q = Q(x + beta * err); err <- (x + beta * err) - DQ(q). err: fp32 [n].
int dsb200_loco_group_quant(const void* x, int dt, float* err, void* q, float* scales, int64_t n, int group, int bits,
                            float beta, int swz_nodes, int swz_devs, cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (group % 256 != 0 || n % group != 0 || (bits != 8 && bits != 4)) {
        dsb200_set_error("loco_group_quant: group must be a multiple of 256 dividing n; bits in {4,8}");
        return -1;
    }
    const int64_t ng = n / group;
    if (swz_nodes * swz_devs > 0 && ng % ((int64_t)swz_nodes * swz_devs) != 0) {
        dsb200_set_error("loco_group_quant: groups must divide evenly over nodes*devices for the swizzled layout");
        return -1;
    }
#define LQ(T, B)                                                                                                      \
    loco_group_quant_kernel<T, B><<<grid_warps(ng), 256, 0, stream>>>((const T*)x, err, (int8_t*)q, scales, ng, group, \
                                                                      beta, swz_nodes, swz_devs)
    if (dt == DSB_DT_BF16) {
        if (bits == 8) LQ(__nv_bfloat16, 8); else LQ(__nv_bfloat16, 4);
    } else if (dt == DSB_DT_F16) {
        if (bits == 8) LQ(__half, 8); else LQ(__half, 4);
    } else {
        dsb200_set_error("loco_group_quant: input must be fp16/bf16");
        return -1;
    }
#undef LQ
    DSB_CHECK_LAUNCH("loco_group_quant");
    return 0;
}

// In-place fake quantization (quantize-dequantize) of x [n] in groups; bits in [2, 16].
int dsb200_fake_quant(void* x, int dt, int64_t n, int group, int bits, int symmetric, int stochastic, unsigned int seed,
                      cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (group % 256 != 0 || n % group != 0 || bits < 2 || bits > 16) {
        dsb200_set_error("fake_quant: group must be a multiple of 256 dividing n; bits in [2,16]");
        return -1;
    }
    const int64_t ng = n / group;
    if (dt == DSB_DT_BF16)
        fake_quant_kernel<__nv_bfloat16><<<grid_warps(ng), 256, 0, stream>>>((__nv_bfloat16*)x, ng, group, bits, symmetric,
                                                                             stochastic, seed);
    else if (dt == DSB_DT_F16)
        fake_quant_kernel<__half><<<grid_warps(ng), 256, 0, stream>>>((__half*)x, ng, group, bits, symmetric, stochastic,
                                                                      seed);
    else {
        dsb200_set_error("fake_quant: input must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("fake_quant");
    return 0;
}

int dsb200_int_group_dequant(const void* q, const float* scales, void* y, int dt, int64_t n, int group, int bits,
                             cudaStream_t stream)
{
    if (n <= 0) return 0;
#define IDQ(T, B) \
    int_group_dequant_kernel<T, B><<<grid_elems(n / 8), 256, 0, stream>>>((const int8_t*)q, scales, (T*)y, n, group)
    if (dt == DSB_DT_BF16) {
        if (bits == 8) IDQ(__nv_bfloat16, 8); else IDQ(__nv_bfloat16, 4);
    } else if (dt == DSB_DT_F16) {
        if (bits == 8) IDQ(__half, 8); else IDQ(__half, 4);
    } else {
        dsb200_set_error("int_group_dequant: output must be fp16/bf16");
        return -1;
    }
#undef IDQ
    DSB_CHECK_LAUNCH("int_group_dequant");
    return 0;
}

int dsb200_dequant_reduce(const void* q, const float* scales, void* out, int dt, int64_t chunk_elems, int nchunks,
                          int group, int bits, float post, cudaStream_t stream)
{
    if (chunk_elems <= 0) return 0;
#define DR(T, B)                                                                                                    \
    dequant_reduce_kernel<T, B><<<grid_elems(chunk_elems / 8), 256, 0, stream>>>((const int8_t*)q, scales, (T*)out, \
                                                                                 chunk_elems, nchunks, group, post)
    if (dt == DSB_DT_BF16) {
        if (bits == 8) DR(__nv_bfloat16, 8); else DR(__nv_bfloat16, 4);
    } else if (dt == DSB_DT_F16) {
        if (bits == 8) DR(__half, 8); else DR(__half, 4);
    } else {
        dsb200_set_error("dequant_reduce: output must be fp16/bf16");
        return -1;
    }
#undef DR
    DSB_CHECK_LAUNCH("dequant_reduce");
    return 0;
}

int dsb200_minifloat_quant(const void* x, void* y, float* scales, int dt, int64_t n, int group, int exp_bits,
                           int man_bits, float fmt_max, int stochastic, unsigned int seed, cudaStream_t stream)
{
    if (n <= 0) return 0;
    if (group % 256 != 0 || n % group != 0) {
        dsb200_set_error("minifloat_quant: group must be a multiple of 256 dividing n");
        return -1;
    }
    const int64_t ng = n / group;
    if (dt == DSB_DT_BF16)
        minifloat_quant_kernel<__nv_bfloat16><<<grid_warps(ng), 256, 0, stream>>>(
            (const __nv_bfloat16*)x, (__nv_bfloat16*)y, scales, ng, group, exp_bits, man_bits, fmt_max, stochastic, seed);
    else if (dt == DSB_DT_F16)
        minifloat_quant_kernel<__half><<<grid_warps(ng), 256, 0, stream>>>((const __half*)x, (__half*)y, scales, ng, group,
                                                                           exp_bits, man_bits, fmt_max, stochastic, seed);
    else {
        dsb200_set_error("minifloat_quant: input must be fp16/bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("minifloat_quant");
    return 0;
}

}  // extern "C"
