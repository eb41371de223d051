"""``matmul_fp8``: activation x FP8-quantised weight (reference ``ops/fp_quantizer/fp8_gemm.py``; its fast path is a
Triton kernel).

``weight`` holds the FP8 codes of a ``[K, N]`` matrix quantised by :class:`FP_Quantize` in groups of
``quantization_group_size`` values, ``scale`` its group scales. When the quantised tensor is stored the way
``csrc/wo_gemm.cu`` reads it (``[N, K]`` rows, groups along ``K``: what ``deepspeed_b200.linear`` /
``inference/quantization`` produce) and the call is small enough to be weight-bandwidth bound (<= 64 tokens), the
weight-only tcgen05 GEMM consumes the codes directly; otherwise the weight is dequantised once and multiplied by the
library GEMM."""
import torch


def matmul_fp8_fallback(inp, weight, scale, quantization_group_size, quantizer):
    return torch.matmul(inp, quantizer.dequantize(weight, scale=scale))


def matmul_fp8(inp, weight, scale, quantization_group_size, quantizer, weight_is_nk=False):
    """``inp [.., K] @ W`` with ``W`` given as FP8 codes. ``weight_is_nk``: the codes are ``[N, K]`` rows (``W^T``)."""
    if weight_is_nk and inp.is_cuda and weight.dtype == torch.uint8 and weight.dim() == 2:
        from .. import kernels as K
        x2 = inp.reshape(-1, inp.shape[-1])
        N, Kd = weight.shape
        if x2.dtype == torch.bfloat16 and K.wo_gemm_supported(x2, weight, quantization_group_size, 8, Kd):
            scales = scale.reshape(N, -1).float().contiguous()
            return K.wo_gemm(x2, weight, scales, N, Kd, quantization_group_size, 8).view(*inp.shape[:-1], N)
    if weight_is_nk:
        return torch.matmul(inp, quantizer.dequantize(weight, scale=scale).t())
    return matmul_fp8_fallback(inp, weight, scale, quantization_group_size, quantizer)
