"""Straight-through quantizers and the top-k binarizer used by compression-aware layers (reference
``compression/utils.py``). All are autograd functions whose backward passes the gradient through unchanged."""
import math

import torch
from torch import autograd


class TopKBinarizer(autograd.Function):
    """mask = 1 for the ``threshold`` fraction of largest entries of ``inputs``; gradient flows to the scores."""

    @staticmethod
    def forward(ctx, inputs, threshold, sigmoid):
        if sigmoid:
            threshold = torch.sigmoid(threshold).item()
        ctx.sigmoid = sigmoid
        keep = math.ceil(threshold * inputs.numel())
        mask = torch.zeros_like(inputs)
        if keep > 0:
            idx = inputs.flatten().topk(keep, sorted=False).indices
            mask.view(-1)[idx] = 1.0
        ctx.save_for_backward(mask)
        return mask

    @staticmethod
    def backward(ctx, grad_output):
        mask, = ctx.saved_tensors
        if ctx.sigmoid:
            return grad_output.clone(), (grad_output * mask).sum().view(-1), None
        return grad_output.clone(), None, None


class _STE(autograd.Function):

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class SymQuantizer(_STE):
    """Group-wise symmetric uniform quantize-dequantize to ``num_bits``."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None) == (max_value is None) and (min_value is None or num_groups == 1)
        q_range = 2**num_bits
        shape = input.shape
        if min_value is None:
            input = input.reshape(num_groups, -1)
            max_input = input.abs().amax(dim=-1, keepdim=True)
        else:
            max_input = torch.max(min_value.abs(), max_value).view(-1)
        scale = (2 * max_input / q_range).clamp_min(1e-12)
        out = (input / scale).round().clamp(-q_range // 2, q_range // 2 - 1) * scale
        return out.reshape(shape).contiguous()


class AsymQuantizer(_STE):
    """Group-wise asymmetric (min/max) uniform quantize-dequantize."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None) == (max_value is None) and (min_value is None or num_groups == 1)
        q_range = 2**num_bits
        shape = input.shape
        if min_value is None:
            input = input.reshape(num_groups, -1)
            min_value = input.amin(dim=-1, keepdim=True)
            max_value = input.amax(dim=-1, keepdim=True)
        scale = ((max_value - min_value) / q_range).clamp_min(1e-12)
        zero_point = (min_value / scale).round() * scale
        out = ((input - zero_point) / scale).round().clamp(0, q_range - 1) * scale + zero_point
        return out.reshape(shape).contiguous()


class TernaryQuantizer(_STE):
    """{-a, 0, +a} per group with threshold 0.7 * mean|w| (TWN)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        flat = input.reshape(num_groups, -1)
        n = flat.shape[1]
        thres = (0.7 * flat.norm(p=1, dim=1) / n).view(-1, 1)
        pos, neg = (flat > thres).type(input.type()), (flat < -thres).type(input.type())
        mask = (flat.abs() > thres).type(input.type())
        alpha = ((mask * flat).abs().sum(dim=1) / mask.sum(dim=1).clamp_min(1)).view(-1, 1)
        return (alpha * pos - alpha * neg).reshape(input.shape).contiguous()


class BinaryQuantizer(_STE):
    """sign(w) * mean|w| per group (XNOR-style)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        flat = input.reshape(num_groups, -1)
        scale = (flat.norm(p=1, dim=1, keepdim=True) / flat.shape[1])
        return flat.sign().mul(scale).reshape(input.shape).contiguous()
