// Channels-last bias-add family for diffusion UNets/VAEs (reference csrc/spatial/csrc/opt_bias_add.cu):
//   out = (x + bias) [+ other [+ other_bias]]     x, other: [batch, seq, channels]; biases: [channels]
// One pass, 16-byte vectors; the bias vector is tiny and stays in L1.
#include "common.cuh"

namespace {

template <typename T, int MODE>  // 0: x+b   1: x+b+other   2: x+b+other+ob
__global__ void __launch_bounds__(256) bias_add_kernel(const T* __restrict__ x, const T* __restrict__ bias,
                                                       const T* __restrict__ other, const T* __restrict__ other_bias,
                                                       T* __restrict__ out, int64_t nvec, int channels)
{
    constexpr int VEC = 16 / sizeof(T);
    const int cvec = channels / VEC;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(i % cvec) * VEC;
        uint4 xr = dsb::ldg_stream(reinterpret_cast<const uint4*>(x) + i);
        uint4 br = *reinterpret_cast<const uint4*>(bias + c);
        uint4 orr = xr, obr = br;
        if (MODE >= 1) orr = dsb::ldg_stream(reinterpret_cast<const uint4*>(other) + i);
        if (MODE == 2) obr = *reinterpret_cast<const uint4*>(other_bias + c);
        const T *xe = reinterpret_cast<const T*>(&xr), *be = reinterpret_cast<const T*>(&br);
        const T *oe = reinterpret_cast<const T*>(&orr), *obe = reinterpret_cast<const T*>(&obr);
        uint4 res;
        T* re = reinterpret_cast<T*>(&res);
#pragma unroll
        for (int k = 0; k < VEC; k++) {
            float v = dsb::to_f<T>(xe[k]) + dsb::to_f<T>(be[k]);
            if (MODE >= 1) v += dsb::to_f<T>(oe[k]);
            if (MODE == 2) v += dsb::to_f<T>(obe[k]);
            re[k] = dsb::from_f<T>(v);
        }
        reinterpret_cast<uint4*>(out)[i] = res;
    }
}

template <typename T>
int launch(const void* x, const void* bias, const void* other, const void* other_bias, void* out, int64_t n,
           int channels, cudaStream_t stream)
{
    constexpr int VEC = 16 / sizeof(T);
    const int64_t nvec = n / VEC;
    int64_t blocks = (nvec + 255) / 256;
    if (blocks > DSB_NUM_SMS * 8) blocks = DSB_NUM_SMS * 8;
    if (blocks < 1) blocks = 1;
    const int mode = other == nullptr ? 0 : (other_bias == nullptr ? 1 : 2);
#define GO(M) bias_add_kernel<T, M><<<(int)blocks, 256, 0, stream>>>((const T*)x, (const T*)bias, (const T*)other, (const T*)other_bias, (T*)out, nvec, channels)
    if (mode == 0) GO(0); else if (mode == 1) GO(1); else GO(2);
#undef GO
    return 0;
}

}  // namespace

// x/other/out: n = batch * seq * channels elements (channels innermost); dt 1 fp16 / 2 bf16; channels % 8 == 0.
extern "C" int dsb200_nhwc_bias_add(const void* x, const void* bias, const void* other, const void* other_bias,
                                    void* out, int64_t n, int channels, int dt, cudaStream_t stream)
{
    if (channels % 8 != 0 || n % channels != 0) {
        dsb200_set_error("nhwc_bias_add: channels must be a multiple of 8 and divide the element count");
        return -1;
    }
    if (dt == DSB_DT_F16) launch<__half>(x, bias, other, other_bias, out, n, channels, stream);
    else if (dt == DSB_DT_BF16) launch<__nv_bfloat16>(x, bias, other, other_bias, out, n, channels, stream);
    else {
        dsb200_set_error("nhwc_bias_add: dtype must be fp16 or bf16");
        return -1;
    }
    DSB_CHECK_LAUNCH("nhwc_bias_add");
    return 0;
}
