// Shared device helpers for the sm_100a kernel library (deepspeed_b200).
// Conventions: dtype codes 0=fp32 1=fp16 2=bf16; all launchers return 0 on success and record the CUDA error
// string otherwise (dsb200_last_error()).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define DSB_DT_F32 0
#define DSB_DT_F16 1
#define DSB_DT_BF16 2

#define DSB_NUM_SMS 148

extern "C" const char* dsb200_last_error();
void dsb200_set_error(const char* msg);

#define DSB_CHECK_LAUNCH(name)                                   \
    do {                                                         \
        cudaError_t _e = cudaGetLastError();                     \
        if (_e != cudaSuccess) {                                 \
            dsb200_set_error(cudaGetErrorString(_e));            \
            return (int)_e;                                      \
        }                                                        \
    } while (0)

#define DSB_CHECK(call)                                          \
    do {                                                         \
        cudaError_t _e = (call);                                 \
        if (_e != cudaSuccess) {                                 \
            dsb200_set_error(cudaGetErrorString(_e));            \
            return (int)_e;                                      \
        }                                                        \
    } while (0)

namespace dsb {

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum; every thread gets the result. `smem` needs 33 floats. Safe to call repeatedly.
__device__ __forceinline__ float block_sum(float v, float* smem)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect smem from a previous call
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? smem[lane] : 0.f;
    r = warp_sum(r);
    return r;
}

__device__ __forceinline__ float block_max(float v, float* smem)
{
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    float r = (lane < nw) ? smem[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

// ---- 8 x 16-bit <-> 8 floats ------------------------------------------------------------------------
struct f8 {
    float v[8];
};

template <typename T>
__device__ __forceinline__ float to_f(T x);
template <>
__device__ __forceinline__ float to_f<float>(float x) { return x; }
template <>
__device__ __forceinline__ float to_f<__half>(__half x) { return __half2float(x); }
template <>
__device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 x) { return __bfloat162float(x); }

template <typename T>
__device__ __forceinline__ T from_f(float x);
template <>
__device__ __forceinline__ float from_f<float>(float x) { return x; }
template <>
__device__ __forceinline__ __half from_f<__half>(float x) { return __float2half_rn(x); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

// 16-byte load of 8 16-bit values
template <typename T>
__device__ __forceinline__ f8 load8(const T* p)
{
    static_assert(sizeof(T) == 2, "load8 is for 16-bit types");
    uint4 raw = *reinterpret_cast<const uint4*>(p);
    const T* h = reinterpret_cast<const T*>(&raw);
    f8 r;
#pragma unroll
    for (int i = 0; i < 8; i++) r.v[i] = to_f<T>(h[i]);
    return r;
}

template <typename T>
__device__ __forceinline__ void store8(T* p, const f8& x)
{
    static_assert(sizeof(T) == 2, "store8 is for 16-bit types");
    uint4 raw;
    T* h = reinterpret_cast<T*>(&raw);
#pragma unroll
    for (int i = 0; i < 8; i++) h[i] = from_f<T>(x.v[i]);
    *reinterpret_cast<uint4*>(p) = raw;
}

// streaming (evict-first) variants for data touched once
__device__ __forceinline__ uint4 ldg_stream(const void* p)
{
    uint4 r;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// generic scalar load/store by dtype code (block-uniform branch)
__device__ __forceinline__ float ld_dt(const void* p, int dt, int64_t i)
{
    if (dt == DSB_DT_F32) return reinterpret_cast<const float*>(p)[i];
    if (dt == DSB_DT_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    return __half2float(reinterpret_cast<const __half*>(p)[i]);
}
__device__ __forceinline__ void st_dt(void* p, int dt, int64_t i, float v)
{
    if (dt == DSB_DT_F32)
        reinterpret_cast<float*>(p)[i] = v;
    else if (dt == DSB_DT_BF16)
        reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
    else
        reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
}

// 4-element vector load by dtype code: fp32 -> one 16 B load, 16-bit -> one 8 B load
__device__ __forceinline__ float4 ld4_dt(const void* p, int dt, int64_t i4)
{
    if (dt == DSB_DT_F32) return reinterpret_cast<const float4*>(p)[i4];
    uint2 raw = reinterpret_cast<const uint2*>(p)[i4];
    float4 r;
    if (dt == DSB_DT_BF16) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
        float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
        r = make_float4(a.x, a.y, b.x, b.y);
    } else {
        const __half2* h = reinterpret_cast<const __half2*>(&raw);
        float2 a = __half22float2(h[0]), b = __half22float2(h[1]);
        r = make_float4(a.x, a.y, b.x, b.y);
    }
    return r;
}
__device__ __forceinline__ void st4_dt(void* p, int dt, int64_t i4, float4 v)
{
    if (dt == DSB_DT_F32) {
        reinterpret_cast<float4*>(p)[i4] = v;
        return;
    }
    uint2 raw;
    if (dt == DSB_DT_BF16) {
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
        h[0] = __floats2bfloat162_rn(v.x, v.y);
        h[1] = __floats2bfloat162_rn(v.z, v.w);
    } else {
        __half2* h = reinterpret_cast<__half2*>(&raw);
        h[0] = __floats2half2_rn(v.x, v.y);
        h[1] = __floats2half2_rn(v.z, v.w);
    }
    reinterpret_cast<uint2*>(p)[i4] = raw;
}

__device__ __forceinline__ bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace dsb
