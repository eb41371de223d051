"""MoQ: quantize weights *during training* with a bit-width schedule (reference ``runtime/quantize.py:14``;
kernels ``csrc/quantization/fake_quantizer.cu``: symmetric/asymmetric, nearest/stochastic rounding per group)."""
import math

import torch

TWO_D_PARAMS = 6


class Quantizer:

    def __init__(self, q_groups=1, q_mixed_fp16=False, q_change_ratio=0.01, q_type=0, q_rounding=0, q_verbose=False,
                 q_eigenvalue=False, use_quantizer_kernel=False, layer_num=0):
        self.q_groups, self.q_mixed_fp16, self.q_change_ratio = q_groups, q_mixed_fp16, q_change_ratio
        self.q_type, self.q_rounding, self.q_verbose, self.q_eigenvalue = q_type, q_rounding, q_verbose, q_eigenvalue
        self.qsteps = 0
        self.quantize_real_ratio = 1.000
        self.layer_num = layer_num
        self.q_start_bits = self.q_target_bits = self.q_period = None

    def any_precision_switch(self):
        if self.layer_num == 0:
            return True
        return any(b != t for b, t in zip(self.q_start_bits, self.q_target_bits))

    def quantize(self, parameter_group, overflow, eigenvalue_enabled, block_eigenvalue=None):
        if overflow and not eigenvalue_enabled:
            return
        self.step()
        self.update_fp16_ratio()
        for group in parameter_group:
            for p in group:
                if not (p.dim() > 1 and hasattr(p, "start_bits") and p.start_bits):
                    continue
                factor = 1.0
                if block_eigenvalue and id(p) in block_eigenvalue:
                    ev, _ = block_eigenvalue[id(p)]
                    factor = 1 + math.floor(ev * 4) if ev is not None else 1.0
                p.data = self.compute_quantization(p, factor=factor)

    def step(self):
        self.qsteps += 1

    def update_fp16_ratio(self):
        if self.q_mixed_fp16:
            self.quantize_real_ratio = max(0.0, self.quantize_real_ratio - self.q_change_ratio)

    @staticmethod
    def fake_quantize(x, bits, groups, symmetric=True, stochastic=False):
        """Quantize-dequantize ``x`` group-wise to ``bits`` bits. Half-precision CUDA tensors go through the in-place
        kernel (``csrc/quant.cu`` ``fake_quant_kernel``: one read and one write of the tensor)."""
        if x.is_cuda and x.dtype in (torch.bfloat16, torch.float16) and x.numel() % groups == 0 and \
                (x.numel() // groups) % 256 == 0:
            from ..ops import quantizer as Q
            if Q._cuda(x) and hasattr(Q._lib(), "dsb200_fake_quant"):
                seed = int(torch.randint(0, 2**31 - 1, (1, )).item()) if stochastic else 0
                return Q.fake_quantize(x.contiguous().clone(), groups, bits, symmetric, stochastic, seed)
        flat = x.reshape(groups, -1).float()
        if symmetric:
            qmax = 2**(bits - 1) - 1
            scale = flat.abs().amax(dim=1, keepdim=True).clamp(min=1e-8) / qmax
            q = flat / scale
            q = torch.floor(q + torch.rand_like(q)) if stochastic else torch.round(q)
            out = q.clamp(-qmax - 1, qmax) * scale
        else:
            lo, hi = flat.amin(dim=1, keepdim=True), flat.amax(dim=1, keepdim=True)
            scale = (hi - lo).clamp(min=1e-8) / (2**bits - 1)
            q = (flat - lo) / scale
            q = torch.floor(q + torch.rand_like(q)) if stochastic else torch.round(q)
            out = q.clamp(0, 2**bits - 1) * scale + lo
        return out.reshape(x.shape).to(x.dtype)

    def compute_quantization(self, p, index=0, factor=1):
        if p.start_bits != p.target_bits and self.qsteps >= p.q_period:
            self.quantize_real_ratio = 1.0
            p.q_period <<= 1
            p.q_period = int(p.q_period * factor)
            p.start_bits -= 1
        assert p.start_bits >= p.target_bits, "Quantization bit is lower than target precision bits!"
        q = self.fake_quantize(p.data, p.start_bits, self.q_groups, symmetric=(self.q_type == 0),
                               stochastic=(self.q_rounding == 1))
        if self.q_mixed_fp16 and p.start_bits >= p.target_bits - 1:
            q = self.quantize_real_ratio * p.data + (1 - self.quantize_real_ratio) * q
        return q
