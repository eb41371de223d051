"""v1 inference transformer layer (reference ``model_implementations/transformers/ds_transformer.py``
``DeepSpeedTransformerInference`` with ``ops/transformer/inference/{ds_attention,ds_mlp}.py``): one decoder layer whose
weights are plain tensors filled by an injection policy, with an internal KV cache for token-by-token generation.

The reference drives ~15 kernels of ``csrc/transformer/inference`` through a shared workspace; here the layer is a
handful of calls: fused norm kernels (``ops/fused.py``), the library GEMM, rotary on the q/k views, and scaled-dot-product
attention over the cached keys / values. Parameter names (``attention.attn_qkvw`` ... ``mlp.output_b``, ``norm_w``) are
the reference's, so containers and checkpoints written for it fill this layer unchanged."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .... import comm as dist
from ....utils.types import GATED_ACTIVATION_TYPES, ActivationFuncType, NormType
from ... import fused


def _empty(cfg, *shape, dtype=None):
    if cfg.set_empty_params:
        return None
    return nn.Parameter(torch.empty(*shape, dtype=dtype or (torch.half if cfg.dtype == torch.int8 else cfg.dtype)),
                        requires_grad=False)


def _linear(x, w, b, transposed):
    if transposed:  # weight stored [in, out]
        y = x @ w
        return y + b if b is not None else y
    return F.linear(x, w, b)


def _norm(cfg, x, w, b):
    if cfg.norm_type == NormType.RMSNorm:
        return fused.rms_norm(x, w, cfg.epsilon)
    return fused.layer_norm(x, w, b, cfg.epsilon)


def _rotary(x, pos, cfg):
    """x: [b, heads, s, d]; rotate the first ``rotary_dim`` channels. ``rotate_half``: NeoX / Llama pairing (i, i + r/2);
    ``rotate_every_two``: GPT-J pairing (2i, 2i + 1)."""
    r = cfg.rotary_dim
    if r <= 0:
        return x
    inv = 1.0 / (float(cfg.rope_theta)**(torch.arange(0, r, 2, device=x.device, dtype=torch.float32) / r))
    ang = pos.to(torch.float32)[:, None] * inv[None, :]  # [s, r/2]
    cos, sin = ang.cos()[None, None], ang.sin()[None, None]
    xr, rest = x[..., :r].float(), x[..., r:]
    if cfg.rotate_half:
        a, b = xr[..., :r // 2], xr[..., r // 2:]
        out = torch.cat([a * cos - b * sin, b * cos + a * sin], dim=-1)
    else:
        a, b = xr[..., 0::2], xr[..., 1::2]
        out = torch.stack([a * cos - b * sin, b * cos + a * sin], dim=-1).flatten(-2)
    return torch.cat([out.to(x.dtype), rest], dim=-1)


class DeepSpeedSelfAttention(nn.Module):
    num_layers = 0

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1):
        super().__init__()
        self.config, self.mp_group = config, mp_group
        DeepSpeedSelfAttention.num_layers += 1
        h, mp = config.hidden_size, config.mp_size
        self.heads = config.heads // mp
        self.kv_heads = (config.num_kv if config.num_kv > 0 else config.heads) // mp
        self.head_dim = h // config.heads
        qkv_out = (self.heads + 2 * self.kv_heads) * self.head_dim
        t = config.transposed_mode
        self.attn_qkvw = _empty(config, *((h, qkv_out) if t else (qkv_out, h)))
        self.attn_qkvb = _empty(config, qkv_out)
        self.attn_ow = _empty(config, *((h // mp, h) if t else (h, h // mp)))
        self.attn_ob = _empty(config, h)
        self.k_cache = self.v_cache = None

    def reset_cache(self):
        self.k_cache = self.v_cache = None

    def forward(self, x, input_mask=None, layer_past=None, use_cache=False, alibi=None):
        cfg = self.config
        b, s, _ = x.shape
        qkv = _linear(x, self.attn_qkvw, self.attn_qkvb, cfg.transposed_mode)
        hq, hk, d = self.heads, self.kv_heads, self.head_dim
        if cfg.bigscience_bloom:  # BLOOM interleaves q, k, v per head: [heads, 3, d]
            qkv = qkv.view(b, s, hq, 3, d)
            q, k, v = qkv[..., 0, :], qkv[..., 1, :], qkv[..., 2, :]
        else:
            q, k, v = qkv.split([hq * d, hk * d, hk * d], dim=-1)
            q, k, v = q.view(b, s, hq, d), k.view(b, s, hk, d), v.view(b, s, hk, d)
        q, k, v = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
        if layer_past is not None:
            past_k, past_v = layer_past[0], layer_past[1]
        else:
            past_k, past_v = self.k_cache, self.v_cache
        past = 0 if past_k is None else past_k.shape[2]
        pos = torch.arange(past, past + s, device=x.device)
        q, k = _rotary(q, pos, cfg), _rotary(k, pos, cfg)
        if past_k is not None:
            k, v = torch.cat([past_k, k], dim=2), torch.cat([past_v, v], dim=2)
        if use_cache or layer_past is None:
            self.k_cache, self.v_cache = k, v
        total = k.shape[2]
        scale = 1.0 / math.sqrt(d) if cfg.scale_attention else 1.0
        if cfg.scale_attn_by_inverse_layer_idx:
            scale = scale / float(cfg.layer_id + 1)
        # additive bias: causal / local window / padding mask / alibi
        bias = None
        if cfg.triangular_masking or cfg.local_attention:
            qi = torch.arange(past, past + s, device=x.device)[:, None]
            ki = torch.arange(total, device=x.device)[None, :]
            ok = torch.ones(s, total, dtype=torch.bool, device=x.device)
            if cfg.triangular_masking:
                ok &= ki <= qi
            if cfg.local_attention:
                ok &= ki > qi - cfg.window_size
            bias = torch.zeros(s, total, dtype=torch.float32, device=x.device).masked_fill(~ok, float("-inf"))[None, None]
        if input_mask is not None:
            m = input_mask
            additive = m.is_floating_point() and bool((m < 0).any())
            if not additive:
                # 0/1 mask: with invert_mask (default) 1 marks a position to attend to, otherwise a masked one
                # (BLOOM hands the already inverted mask)
                keep = m.bool() if cfg.invert_mask else ~m.bool()
                m = torch.zeros(m.shape, dtype=torch.float32, device=x.device).masked_fill(~keep, float("-inf"))
            while m.dim() < 4:
                m = m.unsqueeze(1)
            m = m[..., -total:].float()
            bias = m if bias is None else bias + m
        if alibi is not None:
            a = alibi.float().view(b, -1, 1, alibi.shape[-1])[..., -total:] if alibi.dim() == 3 else alibi.float()
            if a.shape[1] != hq:  # full-head alibi given to a tensor-parallel rank
                r = dist.get_rank(self.mp_group) if self.mp_group is not None else 0
                a = a[:, r * hq:(r + 1) * hq]
            bias = a if bias is None else bias + a
        if hk != hq:
            k = k.repeat_interleave(hq // hk, dim=1)
            v = v.repeat_interleave(hq // hk, dim=1)
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=None if bias is None else bias.to(q.dtype), scale=scale)
        ctx = ctx.transpose(1, 2).reshape(b, s, hq * d)
        out = _linear(ctx, self.attn_ow, None, cfg.transposed_mode)
        if cfg.mp_size > 1 and self.mp_group is not None:
            dist.all_reduce(out, group=self.mp_group)
        return out, (self.k_cache, self.v_cache) if (use_cache or layer_past is None) else (k, v)


class BloomSelfAttention(DeepSpeedSelfAttention):
    pass  # the BLOOM q/k/v interleaving is selected by config.bigscience_bloom


class DeepSpeedMLP(nn.Module):

    def __init__(self, config, mp_group=None, q_scales=None, q_groups=1, merge_count=1, mlp_extra_grouping=False):
        super().__init__()
        self.config, self.mp_group = config, mp_group
        h, inter, mp, t = config.hidden_size, config.intermediate_size, config.mp_size, config.transposed_mode
        gated = config.mlp_act_func_type in GATED_ACTIVATION_TYPES
        fan = (2 if gated else 1) * inter // mp
        self.attn_nw = _empty(config, h)  # post-attention norm (pre-MLP), reference naming
        self.attn_nb = _empty(config, h)
        self.inter_w = _empty(config, *((h, fan) if t else (fan, h)))
        self.inter_b = _empty(config, fan)
        self.output_w = _empty(config, *((inter // mp, h) if t else (h, inter // mp)))
        self.output_b = _empty(config, h)

    def _act(self, y):
        a = self.config.mlp_act_func_type
        if a == ActivationFuncType.ReLU:
            return F.relu(y)
        if a == ActivationFuncType.GATED_GELU:
            up, gate = y.chunk(2, dim=-1)
            return up * F.gelu(gate, approximate="tanh")
        if a == ActivationFuncType.GATED_SILU:
            up, gate = y.chunk(2, dim=-1)
            return up * F.silu(gate)
        return F.gelu(y, approximate="tanh")

    def forward(self, x):
        cfg = self.config
        y = self._act(_linear(x, self.inter_w, self.inter_b, cfg.transposed_mode))
        out = _linear(y, self.output_w, None, cfg.transposed_mode)
        if cfg.mp_size > 1 and self.mp_group is not None:
            dist.all_reduce(out, group=self.mp_group)
        return out


class DeepSpeedTransformerInference(nn.Module):
    layer_id = 0

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1,
                 mlp_extra_grouping=False):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerInference.layer_id
        DeepSpeedTransformerInference.layer_id += 1
        attn_cls = BloomSelfAttention if config.bigscience_bloom else DeepSpeedSelfAttention
        self.attention = attn_cls(config, mp_group, quantize_scales, quantize_groups, merge_count)
        self.mlp = DeepSpeedMLP(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
        self.norm_w = _empty(config, config.hidden_size)  # input norm (pre-LN) / final norm (post-LN)
        self.norm_b = _empty(config, config.hidden_size)
        self.layer_past = None

    @classmethod
    def reset_cache(cls, model=None):
        """Drop the internal KV caches (start of a new generation). The reference resets one global workspace; caches
        are per layer here, so pass the model (or call ``layer.attention.reset_cache()``)."""
        if model is not None:
            for m in model.modules():
                if isinstance(m, DeepSpeedSelfAttention):
                    m.reset_cache()

    def forward(self, input=None, input_mask=None, attention_mask=None, attn_mask=None, head_mask=None,
                layer_past=None, get_key_value=False, get_present=False, encoder_output=None, enc_dec_attn_mask=None,
                x=None, encoder_hidden_states=None, encoder_attention_mask=None, use_cache=False, alibi=None,
                output_attentions=False, layer_head_mask=None, past_key_value=None, **kwargs):
        if x is not None:
            input = x
        if "hidden_states" in kwargs:
            input = kwargs["hidden_states"]
        if layer_past is not None and past_key_value is not None:
            raise ValueError("Only one of `layer_past` or `past_key_value` can be present.")
        layer_past = layer_past if layer_past is not None else past_key_value
        mask = attention_mask if attention_mask is not None else (attn_mask if attn_mask is not None else input_mask)
        get_present = get_present or get_key_value or use_cache
        cfg = self.config
        squeeze = input.dim() == 2
        h = input.unsqueeze(0) if squeeze else input
        in_dtype = h.dtype
        wdtype = self.attention.attn_qkvw.dtype
        h = h.to(wdtype)
        if cfg.pre_layer_norm:
            a_in = _norm(cfg, h, self.norm_w, self.norm_b)
            attn_out, present = self.attention(a_in, mask, layer_past, get_present, alibi)
            attn_out = attn_out + self.attention.attn_ob
            if cfg.mlp_after_attn:
                h = h + attn_out
                out = h + self.mlp(_norm(cfg, h, self.mlp.attn_nw, self.mlp.attn_nb)) + self.mlp.output_b
            else:  # GPT-J / NeoX: attention and MLP read the same normed input (NeoX: its own second norm)
                m_in = a_in if self.mlp.attn_nw is None or cfg.norm_type is None else \
                    (_norm(cfg, h, self.mlp.attn_nw, self.mlp.attn_nb) if getattr(cfg, "use_second_norm", False)
                     else a_in)
                out = h + attn_out + self.mlp(m_in) + self.mlp.output_b
        else:
            attn_out, present = self.attention(h, mask, layer_past, get_present, alibi)
            h = _norm(cfg, h + attn_out + self.attention.attn_ob, self.mlp.attn_nw, self.mlp.attn_nb)
            out = _norm(cfg, h + self.mlp(h) + self.mlp.output_b, self.norm_w, self.norm_b)
        out = out.to(in_dtype)
        if squeeze:
            out = out.squeeze(0)
        if get_present:
            return (out, present)
        if cfg.return_single_tuple:
            return (out, )
        if cfg.return_tuple:
            return out if isinstance(out, tuple) else (out, mask)
        return out
