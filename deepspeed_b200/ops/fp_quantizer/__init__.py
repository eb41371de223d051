"""FP quantizer (reference ``ops/fp_quantizer/quantize.py`` ``FP_Quantize`` over ``csrc/fp_quantizer``): group-wise
minifloat quantization, q_bits in {8: e4m3, 6: e3m2, 4: e2m1, 12: e4m7}, per-group fp32 scale, optional stochastic
rounding, ``dequantize`` / ``selective_dequantize`` / ``get_scales``.

Storage: 8-bit -> e4m3 bytes (hardware cvt); 4-bit -> two e2m1 codes per byte; 6/12-bit -> the rounded value kept in
bf16 (bit-packing of 6/12-bit payloads is a storage-only optimisation that the kernels do not need).
"""
import ctypes

import torch

from .. import kernels as K
from ..quantizer import _lib

_FORMATS = {8: (4, 3, 448.0), 6: (3, 2, 28.0), 4: (2, 1, 6.0), 12: (4, 7, 510.0)}
_E2M1_VALUES = torch.tensor([0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0])


class Quantizer:

    def __init__(self, group_size=512):
        self.group_size = group_size


class FP_Quantize(Quantizer):

    def __init__(self, quantization_config=None, group_size=512):
        if quantization_config is not None:
            group_size = getattr(quantization_config, "group_size", group_size)
        super().__init__(group_size)
        self.orig_dtype = None
        self.orig_shape = None
        self.num_groups = None
        self.scale = None
        self.q_bits = None

    def _round(self, x, q_bits, stochastic):
        eb, mb, fmax = _FORMATS[q_bits]
        flat = x.contiguous().view(-1)
        n = flat.numel()
        g = self.group_size
        assert n % g == 0, f"numel {n} must be divisible by group_size {g}"
        if flat.is_cuda and K.available() and g % 256 == 0 and flat.dtype in (torch.bfloat16, torch.float16):
            y = torch.empty_like(flat)
            s = torch.empty(n // g, dtype=torch.float32, device=flat.device)
            seed = int(torch.randint(0, 2**31 - 1, (1, )).item()) if stochastic else 0
            K._check(_lib().dsb200_minifloat_quant(K._ptr(flat), K._ptr(y), K._ptr(s), K._dt(flat), n, g, eb, mb, fmax,
                                                   int(stochastic), seed, K._stream()), "minifloat_quant")
            return y.float(), s
        xg = flat.float().view(-1, g)
        s = xg.abs().amax(dim=1) / fmax
        s = torch.where(s > 0, s, torch.ones_like(s))
        v = xg / s[:, None]
        a = v.abs().clamp(max=fmax)
        e = torch.floor(torch.log2(a.clamp(min=1e-30)))
        bias = (1 << (eb - 1)) - 1
        e = e.clamp(min=1 - bias)
        ulp = torch.pow(2.0, e - mb)
        r = torch.rand_like(a) if stochastic else 0.5
        qa = (torch.floor(a / ulp + r) if stochastic else torch.round(a / ulp)) * ulp
        qa = qa.clamp(max=fmax)
        return (torch.sign(v) * qa).view(-1), s

    def quantize(self, input, q_bits=8, q_mantisa_bits=3, stochastic_mode=False, return_meta_tensor=False):
        assert q_bits in _FORMATS, f"q_bits must be one of {sorted(_FORMATS)}"
        self.orig_dtype, self.orig_shape, self.q_bits = input.dtype, input.shape, q_bits
        vals, scale = self._round(input, q_bits, stochastic_mode)
        self.num_groups = scale.numel()
        self.scale = scale
        if q_bits == 8:
            payload = vals.to(torch.float8_e4m3fn).view(torch.uint8)
        elif q_bits == 4:
            table = _E2M1_VALUES.to(vals.device)
            idx = torch.bucketize(vals.abs(), (table[:-1] + table[1:]) / 2).to(torch.uint8)
            code = idx | ((vals < 0).to(torch.uint8) << 3)
            payload = code[0::2] | (code[1::2] << 4)
        else:
            payload = vals.to(torch.bfloat16)
        if return_meta_tensor:
            return payload, scale
        return _QTensor(payload, scale, q_bits, self.group_size, input.shape, input.dtype)

    def _decode(self, payload, q_bits):
        if q_bits == 8:
            return payload.view(torch.float8_e4m3fn).float()
        if q_bits == 4:
            table = _E2M1_VALUES.to(payload.device)
            lo, hi = payload & 0xF, (payload >> 4) & 0xF
            codes = torch.stack([lo, hi], dim=1).view(-1)
            vals = table[(codes & 7).long()]
            return torch.where((codes & 8) > 0, -vals, vals)
        return payload.float()

    def dequantize(self, input_q, fp_out=None, q_bits=8, q_mantisa_bits=3, scale=None):
        if isinstance(input_q, _QTensor):
            payload, scale, q_bits = input_q.payload, input_q.scale, input_q.q_bits
            shape, dtype, g = input_q.shape, input_q.dtype, input_q.group_size
        else:
            payload, shape, dtype, g = input_q, self.orig_shape, self.orig_dtype, self.group_size
            scale = scale if scale is not None else self.scale
        vals = self._decode(payload, q_bits).view(-1, g) * scale.view(-1, 1)
        out = vals.view(shape).to(dtype)
        if fp_out is not None:
            fp_out.copy_(out)
            return fp_out
        return out

    def selective_dequantize(self, input_q, indices, fp_out=None, q_bits=8, q_mantisa_bits=3, scale=None):
        """Dequantize only the rows ``indices`` of the first dimension (expert selection in MoE inference)."""
        full_shape = input_q.shape if isinstance(input_q, _QTensor) else self.orig_shape
        assert len(full_shape) == 3, "Selective-Dequantization works on 3d tensor only! Please reshape the tensor."
        rows = self.dequantize(input_q, q_bits=q_bits, scale=scale)
        out = rows.index_select(0, indices.to(rows.device).long())
        if fp_out is not None:
            fp_out.copy_(out)
            return fp_out
        return out

    def get_scales(self):
        return self.scale


class _QTensor:
    """Quantized payload + metadata (what the reference appends inline after each group's bytes)."""

    def __init__(self, payload, scale, q_bits, group_size, shape, dtype):
        self.payload, self.scale, self.q_bits, self.group_size, self.shape, self.dtype = \
            payload, scale, q_bits, group_size, shape, dtype

    def numel(self):
        return int(torch.Size(self.shape).numel())

    def nbytes(self):
        return self.payload.numel() * self.payload.element_size() + self.scale.numel() * 4


class FPQuantizerBuilder:
    NAME = "fp_quantizer"

    def is_compatible(self, verbose=False):
        return True

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
