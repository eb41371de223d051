"""Channels-last bias-add ops for diffusion models (reference ``csrc/spatial``, ``op_builder/spatial_inference.py``):
``nhwc_bias_add(x, bias)``, ``nhwc_bias_add_add(x, bias, other)``, ``nhwc_bias_add_bias_add(x, bias, other, other_bias)``.
4-D inputs must be channels-last, 3-D inputs are [batch, seq, channels]."""
import ctypes

import torch

from .. import kernels as K


def _channels(x):
    if x.dim() == 4:
        assert x.is_contiguous(memory_format=torch.channels_last), "4-D input must be channels-last"
        return x.size(1)
    assert x.is_contiguous()
    return x.size(-1)


def _run(x, bias, other=None, other_bias=None):
    c = _channels(x)
    if x.is_cuda and K.available() and x.dtype in (torch.float16, torch.bfloat16) and c % 8 == 0:
        out = torch.empty_like(x)
        L = K.lib()
        L.dsb200_nhwc_bias_add.argtypes = [ctypes.c_void_p] * 5 + [ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                                                                 ctypes.c_void_p]
        L.dsb200_nhwc_bias_add.restype = ctypes.c_int
        null = ctypes.c_void_p(0)
        K._check(L.dsb200_nhwc_bias_add(K._ptr(x), K._ptr(bias), K._ptr(other) if other is not None else null,
                                        K._ptr(other_bias) if other_bias is not None else null, K._ptr(out),
                                        x.numel(), c, K._dt(x), K._stream()), "nhwc_bias_add")
        return out
    shape = (1, c, 1, 1) if x.dim() == 4 else (c, )
    out = x + bias.view(shape)
    if other is not None:
        out = out + other
    if other_bias is not None:
        out = out + other_bias.view(shape)
    return out


def nhwc_bias_add(activation, bias):
    return _run(activation, bias)


def nhwc_bias_add_add(activation, bias, other):
    return _run(activation, bias, other)


def nhwc_bias_add_bias_add(activation, bias, other, other_bias):
    return _run(activation, bias, other, other_bias)


class SpatialInferenceBuilder:
    NAME = "spatial_inference"

    def is_compatible(self, verbose=False):
        return K.available()

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
