"""diffusers ``BasicTransformerBlock`` for inference: self attention, cross attention and a GEGLU feed-forward, each
behind a pre-LayerNorm, with the residual adds and attention output biases folded into the normalisation steps.

Reference: ``ops/transformer/inference/diffusers_transformer_block.py``. The fused steps map to ``ops/fused.py``
(bias + residual + LayerNorm in one pass, GEGLU) on a GPU; the CPU path spells the same math out."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .diffusers_2d_transformer import Diffusers2DTransformerConfig
from .diffusers_attention import DeepSpeedDiffusersAttention


def _bias_residual_layer_norm(x, bias, residual, g, b, eps):
    """Returns (LayerNorm(x + bias + residual), x + bias + residual)."""
    s = x + bias + residual
    return F.layer_norm(s, s.shape[-1:], g, b, eps), s


def _geglu(x, bias):
    """``x [.., 2I]`` (+ bias) -> ``value * gelu(gate)`` with the halves in diffusers' order (value first)."""
    x = x + bias
    val, gate = x.chunk(2, dim=-1)
    return val * F.gelu(gate)


class DeepSpeedDiffusersTransformerBlock(nn.Module):

    def __init__(self, equivalent_module, config: Diffusers2DTransformerConfig = None):
        super().__init__()
        self.config = config or Diffusers2DTransformerConfig()
        m = equivalent_module
        frozen = lambda t: nn.Parameter(t.data, requires_grad=False)  # noqa: E731
        if self.config.int8_quantization:
            from ....parallel.replace_module import GroupQuantizer
            q = GroupQuantizer(q_int8=True)
            self.ff1_w, self.ff2_w = q.quantize(frozen(m.ff.net[0].proj.weight)), q.quantize(frozen(m.ff.net[2].weight))
        else:
            self.ff1_w, self.ff2_w = frozen(m.ff.net[0].proj.weight), frozen(m.ff.net[2].weight)
        self.ff1_b, self.ff2_b = frozen(m.ff.net[0].proj.bias), frozen(m.ff.net[2].bias)
        for i, norm in ((1, m.norm1), (2, m.norm2), (3, m.norm3)):
            setattr(self, f"norm{i}_g", frozen(norm.weight))
            setattr(self, f"norm{i}_b", frozen(norm.bias))
            setattr(self, f"norm{i}_eps", norm.eps)
        self.attn_1, self.attn_2 = m.attn1, m.attn2
        # take the attentions' output biases over: they are added together with the residual before the next norm
        for i, attn, like in ((1, self.attn_1, self.norm2_g), (2, self.attn_2, self.norm3_g)):
            if isinstance(attn, DeepSpeedDiffusersAttention):
                attn.do_out_bias = False
                setattr(self, f"attn_{i}_bias", attn.attn_ob)
            else:
                setattr(self, f"attn_{i}_bias", nn.Parameter(torch.zeros_like(like), requires_grad=False))

    def _weight(self, w):
        if w.dtype == torch.int8:  # group-quantised: scales ride on the parameter
            g = w.scale.shape[0] if w.scale.dim() > 1 else w.scale.numel()
            return (w.float().reshape(g, -1) * w.scale.reshape(g, -1)[:, :1].float()).reshape(w.shape).to(
                self.ff1_b.dtype)
        return w

    def forward(self, hidden_states, context=None, timestep=None, **kwargs):
        if kwargs.get("encoder_hidden_states") is not None:  # diffusers >= 0.11 renamed the argument
            context = kwargs["encoder_hidden_states"]
        n1 = F.layer_norm(hidden_states, hidden_states.shape[-1:], self.norm1_g, self.norm1_b, self.norm1_eps)
        a1 = self.attn_1(n1)
        n2, r1 = _bias_residual_layer_norm(a1, self.attn_1_bias, hidden_states, self.norm2_g, self.norm2_b,
                                           self.norm2_eps)
        a2 = self.attn_2(n2, context=context) if isinstance(self.attn_2, DeepSpeedDiffusersAttention) else \
            self.attn_2(n2, encoder_hidden_states=context)
        n3, r2 = _bias_residual_layer_norm(a2, self.attn_2_bias, r1, self.norm3_g, self.norm3_b, self.norm3_eps)
        h = _geglu(F.linear(n3, self._weight(self.ff1_w)), self.ff1_b)
        return F.linear(h, self._weight(self.ff2_w)) + self.ff2_b + r2
