"""FP quantizer (reference ``ops/fp_quantizer/quantize.py`` ``FP_Quantize`` over ``csrc/fp_quantizer``): group-wise
minifloat quantization, q_bits in {8: e4m3, 6: e3m2, 4: e2m1, 12: e4m7}, per-group fp32 scale, optional stochastic
rounding, ``dequantize`` / ``selective_dequantize`` / ``get_scales``.

Storage: 8-bit -> e4m3 bytes (hardware cvt); 4-bit -> two e2m1 codes per byte; 6-bit -> four e3m2 codes in three bytes
(little-endian bit stream: code i occupies bits [6i, 6i+6), the layout ``csrc/wo_gemm.cu`` unpacks in shared memory);
12-bit -> two e4m7 codes in three bytes. The reference appends each group's fp32 scale to the group's bytes; here
the scales are a separate tensor.
"""
import ctypes

import torch

from .. import kernels as K
from ..quantizer import _lib

_FORMATS = {8: (4, 3, 448.0), 6: (3, 2, 28.0), 4: (2, 1, 6.0), 12: (4, 7, 510.0)}
_E2M1_VALUES = torch.tensor([0.0, 0.5, 1.0, 1.5, 2.0, 3.0, 4.0, 6.0])


def encode_minifloat(vals, eb, mb):
    """Exactly representable values -> sign/exponent/mantissa codes (uint8 for <= 8 bits, int16 otherwise).
    No inf/nan encodings (e3m2 / e2m1 / e4m7 use every exponent for finite values)."""
    bias = (1 << (eb - 1)) - 1
    a = vals.float().abs()
    _, ex = torch.frexp(a)  # a = m * 2^ex, m in [0.5, 1)
    e_unb = (ex - 1).clamp(min=1 - bias)
    e_unb = torch.where(a > 0, e_unb, torch.full_like(e_unb, 1 - bias))
    frac = a * torch.pow(2.0, -e_unb.float())  # [0, 2): >= 1 normal, < 1 subnormal
    normal = frac >= 1.0
    e_field = torch.where(normal, e_unb + bias, torch.zeros_like(e_unb))
    m_field = torch.round(torch.where(normal, frac - 1.0, frac) * (1 << mb)).to(torch.int32)
    code = (e_field.to(torch.int32) << mb) | m_field
    code = code | ((vals < 0).to(torch.int32) << (eb + mb))
    return code.to(torch.uint8 if eb + mb + 1 <= 8 else torch.int16)


def decode_minifloat(code, eb, mb):
    bias = (1 << (eb - 1)) - 1
    c = code.to(torch.int32)
    m = (c & ((1 << mb) - 1)).float() / (1 << mb)
    e = (c >> mb) & ((1 << eb) - 1)
    mag = torch.where(e > 0, (1.0 + m) * torch.pow(2.0, (e - bias).float()), m * (2.0**(1 - bias)))
    return torch.where(((c >> (eb + mb)) & 1) > 0, -mag, mag)


def pack_6bit(code):
    """uint8 codes (6 significant bits), numel % 4 == 0 -> 3 bytes per 4 codes."""
    c = code.view(-1, 4).to(torch.int32)
    w = c[:, 0] | (c[:, 1] << 6) | (c[:, 2] << 12) | (c[:, 3] << 18)
    return torch.stack([w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF], dim=1).to(torch.uint8).view(-1)


def unpack_6bit(packed):
    b = packed.view(-1, 3).to(torch.int32)
    w = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
    return torch.stack([w & 0x3F, (w >> 6) & 0x3F, (w >> 12) & 0x3F, (w >> 18) & 0x3F], dim=1).to(torch.uint8).view(-1)


def pack_12bit(code):
    c = code.view(-1, 2).to(torch.int32) & 0xFFF
    w = c[:, 0] | (c[:, 1] << 12)
    return torch.stack([w & 0xFF, (w >> 8) & 0xFF, (w >> 16) & 0xFF], dim=1).to(torch.uint8).view(-1)


def unpack_12bit(packed):
    b = packed.view(-1, 3).to(torch.int32)
    w = b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16)
    return torch.stack([w & 0xFFF, (w >> 12) & 0xFFF], dim=1).to(torch.int16).view(-1)


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
        elif q_bits == 6:
            assert vals.numel() % 4 == 0, "6-bit packing needs numel % 4 == 0"
            payload = pack_6bit(encode_minifloat(vals, 3, 2))
        else:
            assert vals.numel() % 2 == 0, "12-bit packing needs an even numel"
            payload = pack_12bit(encode_minifloat(vals, 4, 7))
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
        if payload.dtype != torch.uint8:  # payloads of older checkpoints: rounded values kept in bf16
            return payload.float()
        if q_bits == 6:
            return decode_minifloat(unpack_6bit(payload), 3, 2)
        return decode_minifloat(unpack_12bit(payload), 4, 7)

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
