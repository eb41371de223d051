"""Self / cross attention of diffusers UNet transformer blocks as one inference module.

Reference: ``ops/transformer/inference/diffusers_attention.py`` (``DeepSpeedDiffusersAttention``: fused QKV GEMM +
Triton flash attention for self attention, three GEMMs + softmax for cross attention, output projection whose bias the
enclosing block may take over). Weights are stored ``[in, out]`` (``x @ w``) like the reference so that containers
written for it load unchanged. The attention itself goes to the flash kernel of ``ops/attention.py`` when the head
size is the 128 it is built for, otherwise to the library's fused SDPA (Stable Diffusion heads are 40 / 80 / 160 wide
and cross attention reads 77 text tokens - shapes a training-oriented tile does not fit)."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ....utils.logging import log_dist


def _attend(q, k, v, scale):
    """q [B, H, Lq, D], k/v [B, H, Lk, D] -> [B, H, Lq, D]."""
    if q.is_cuda and q.shape[-1] == 128 and q.shape[2] == k.shape[2] and q.dtype in (torch.bfloat16, torch.float16):
        from ... import attention as A
        if getattr(A, "flash_supported", lambda *a, **kw: False)(q.transpose(1, 2)):
            return A.flash_attention(q.transpose(1, 2).contiguous(), k.transpose(1, 2).contiguous(),
                                     v.transpose(1, 2).contiguous(), causal=False, scale=scale).transpose(1, 2)
    return F.scaled_dot_product_attention(q, k, v, scale=scale)


class DeepSpeedDiffusersAttentionFunction(torch.autograd.Function):
    """Inference only: the forward of :class:`DeepSpeedDiffusersAttention`; asking for gradients is an error."""

    @staticmethod
    def forward(ctx, input, context, input_mask, module):
        return module._forward(input, context, input_mask)

    @staticmethod
    def backward(ctx, *grads):
        raise RuntimeError("You are running with DeepSpeed Inference mode. Please switch to Training mode for "
                           "running backward!")


class DeepSpeedDiffusersAttention(nn.Module):
    layer_id = 0

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedDiffusersAttention.layer_id
        DeepSpeedDiffusersAttention.layer_id += 1
        if DeepSpeedDiffusersAttention.layer_id == 1:
            log_dist(f"DeepSpeed-Attention config: {self.config.__dict__}", [0])
        h, mp = config.hidden_size, max(1, getattr(config, "mp_size", 1))
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        mk = lambda *shape: nn.Parameter(torch.empty(*shape, dtype=dt), requires_grad=False)  # noqa: E731
        self.attn_qkvw = mk(h, h // mp * 3)
        self.attn_qw, self.attn_kw, self.attn_vw = mk(h, h), mk(h, h), mk(h, h)
        self.attn_qkvb = None  # diffusers attention projections carry no bias; set a tensor to add one
        self.attn_ow = mk(h // mp, h)
        self.attn_ob = mk(h)
        self.do_out_bias = True
        self.num_attention_heads_per_partition = config.heads // mp
        self.hidden_size_per_partition = h // mp
        self.hidden_size_per_attention_head = h // config.heads
        self.norm_factor = math.sqrt(math.sqrt(h // config.heads))
        if getattr(config, "scale_attn_by_inverse_layer_idx", False):
            self.norm_factor *= math.sqrt(self.config.layer_id + 1)

    @classmethod
    def from_module(cls, attn, config):
        """Build from a diffusers ``Attention`` / ``CrossAttention`` (``to_q``, ``to_k``, ``to_v``, ``to_out[0]``)."""
        new = cls(config)
        dt = new.attn_ow.dtype
        q, k, v, o = attn.to_q.weight.data, attn.to_k.weight.data, attn.to_v.weight.data, attn.to_out[0].weight.data
        if k.shape[1] == q.shape[1]:  # self attention: one fused projection
            new.attn_qkvw.data = torch.cat([q, k, v], dim=0).t().contiguous().to(dt)
        else:
            new.attn_qkvw = None
        new.attn_qw.data = q.t().contiguous().to(dt)
        new.attn_kw = nn.Parameter(k.t().contiguous().to(dt), requires_grad=False)
        new.attn_vw = nn.Parameter(v.t().contiguous().to(dt), requires_grad=False)
        new.attn_ow.data = o.t().contiguous().to(dt)
        ob = attn.to_out[0].bias
        new.attn_ob.data = ob.data.to(dt) if ob is not None else torch.zeros_like(new.attn_ob)
        return new.to(q.device)

    def _heads(self, x):
        B, L, _ = x.shape
        return x.view(B, L, self.num_attention_heads_per_partition, -1).transpose(1, 2)

    def _forward(self, input, context=None, input_mask=None):
        dt = self.attn_ow.dtype
        if input.dtype != dt:
            input = input.to(dt)
        scale = 1.0 / (self.norm_factor**2)  # 1 / norm_factor on q and on k: 1 / sqrt(head size)
        if context is None and self.attn_qkvw is not None:
            qkv = torch.matmul(input, self.attn_qkvw)
            if self.attn_qkvb is not None:
                qkv = qkv + self.attn_qkvb
            q, k, v = qkv.chunk(3, dim=-1)
        else:
            ctx = input if context is None else context.to(dt)
            q, k, v = torch.matmul(input, self.attn_qw), torch.matmul(ctx, self.attn_kw), torch.matmul(ctx, self.attn_vw)
        out = _attend(self._heads(q), self._heads(k), self._heads(v), scale)
        out = out.transpose(1, 2).reshape(input.shape[0], input.shape[1], self.hidden_size_per_partition)
        out = torch.matmul(out, self.attn_ow)
        return out + self.attn_ob if self.do_out_bias else out

    def forward(self, input, context=None, input_mask=None, encoder_hidden_states=None, **_):
        if encoder_hidden_states is not None:
            context = encoder_hidden_states
        if torch.is_grad_enabled() and input.requires_grad:
            return DeepSpeedDiffusersAttentionFunction.apply(input, context, input_mask, self)
        return self._forward(input, context, input_mask)
