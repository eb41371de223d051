"""Llama-family decoder (Llama-2/3, Mistral-style GQA) built on the fused sm_100a ops.

This is the flagship training model of the framework (``BASELINE.json``: Llama-3-8B ZeRO-3). Layout decisions:

* ``qkv_proj`` and ``gate_up_proj`` are single fused GEMMs (activations are read once, SwiGLU consumes the fused
  output in place); :func:`hf_to_fused_state_dict` / :func:`fused_to_hf_state_dict` convert to/from the Hugging Face
  parameter names.
* pre-norm residual stream uses the fused ``add + RMSNorm`` kernel (one read, two writes).
* the LM head + loss is the chunked fused linear/cross-entropy (no ``[tokens, vocab]`` fp32 tensor).
* every decoder layer is one ZeRO-3 unit and one activation-checkpoint region.
"""
import math
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import fused


@dataclass
class LlamaConfig:
    vocab_size: int = 128256
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    max_position_embeddings: int = 8192
    rms_norm_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_scaling: Optional[dict] = None
    tie_word_embeddings: bool = False
    gradient_checkpointing: bool = True
    loss_chunk_tokens: int = 4096
    attn_impl: str = "auto"  # auto | sdpa | native
    init_std: float = 0.02
    dtype: torch.dtype = torch.bfloat16

    @property
    def head_dim(self):
        return self.hidden_size // self.num_attention_heads

    @staticmethod
    def llama3_8b(**kw):
        return LlamaConfig(**kw)

    @staticmethod
    def tiny(**kw):
        base = dict(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2,
                    num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=256)
        base.update(kw)
        return LlamaConfig(**base)

    def num_params(self):
        h, i, v, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        kv = self.num_key_value_heads * self.head_dim
        per_layer = h * (h + 2 * kv) + h * h + 3 * h * i + 2 * h
        return L * per_layer + v * h * (1 if self.tie_word_embeddings else 2) + h

    def flops_per_token(self, seq_len, with_recompute=False):
        """Training FLOPs per token (fwd+bwd = 3x fwd matmul FLOPs; causal attention counted at half)."""
        h, i, v, L = self.hidden_size, self.intermediate_size, self.vocab_size, self.num_hidden_layers
        kv = self.num_key_value_heads * self.head_dim
        mm = L * (2 * h * (h + 2 * kv) + 2 * h * h + 6 * h * i) + 2 * h * v
        attn = L * 2 * 2 * seq_len * h * 0.5
        fwd = mm + attn
        return fwd * (4 if with_recompute else 3)


class RMSNorm(nn.Module):

    def __init__(self, dim, eps=1e-5, dtype=None, device=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim, dtype=dtype, device=device))
        self.eps = eps

    def forward(self, x, residual=None):
        return fused.rms_norm(x, self.weight, self.eps, residual)


def attention(q, k, v, causal=True, impl="auto"):
    """q: [B,S,Hq,D], k/v: [B,S,Hkv,D] -> [B,S,Hq,D]. GQA handled without materialising repeated KV."""
    if impl in ("auto", "native") and q.is_cuda:
        from ..ops import attention as native_attn
        if (impl == "native" or native_attn.enabled_by_default()) and native_attn.supported(q, k, v):
            return native_attn.flash_attention(q, k, v, causal=causal)
    qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    gqa = qt.shape[1] != kt.shape[1]
    if gqa and not q.is_cuda:
        rep = qt.shape[1] // kt.shape[1]
        kt, vt, gqa = kt.repeat_interleave(rep, dim=1), vt.repeat_interleave(rep, dim=1), False
    backend = _sdpa_backend() if q.is_cuda else None
    if backend is not None:
        with torch.nn.attention.sdpa_kernel(backend):
            out = F.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, enable_gqa=gqa)
    else:
        out = F.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, enable_gqa=gqa)
    return out.transpose(1, 2)


_SDPA_BACKEND = "unset"


def _sdpa_backend():
    """Library attention backend used until the native tcgen05 kernel serves the shape (env DSB200_SDPA)."""
    global _SDPA_BACKEND
    if _SDPA_BACKEND == "unset":
        import os
        from torch.nn.attention import SDPBackend
        name = os.environ.get("DSB200_SDPA", "auto").lower()
        table = {"cudnn": [SDPBackend.CUDNN_ATTENTION], "flash": [SDPBackend.FLASH_ATTENTION],
                 "efficient": [SDPBackend.EFFICIENT_ATTENTION], "math": [SDPBackend.MATH]}
        _SDPA_BACKEND = table.get(name)
    return _SDPA_BACKEND


class LlamaAttention(nn.Module):

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.cfg = cfg
        self.nh, self.nkv, self.hd = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.head_dim
        qkv_out = (self.nh + 2 * self.nkv) * self.hd
        self.qkv_proj = fused.Linear(cfg.hidden_size, qkv_out, bias=False, dtype=cfg.dtype, device=device)
        self.o_proj = fused.Linear(self.nh * self.hd, cfg.hidden_size, bias=False, dtype=cfg.dtype, device=device)

    def forward(self, x, cos, sin, position_ids=None, sp_attention=None):
        B, S, _ = x.shape
        qkv = self.qkv_proj(x)
        # RoPE in place on the q and k heads of the fused projection; q / k / v stay views of it
        q, k, v = fused.apply_rope_qkv(qkv, cos, sin, position_ids, self.nh, self.nkv, self.hd)
        if sp_attention is not None:
            o = sp_attention(q.contiguous(), k.contiguous(), v.contiguous())
        else:
            o = attention(q, k, v, causal=True, impl=self.cfg.attn_impl)
        return self.o_proj(o.reshape(B, S, self.nh * self.hd))


class LlamaMLP(nn.Module):

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.gate_up_proj = fused.Linear(cfg.hidden_size, 2 * cfg.intermediate_size, bias=False, dtype=cfg.dtype,
                                         device=device)
        self.down_proj = fused.Linear(cfg.intermediate_size, cfg.hidden_size, bias=False, dtype=cfg.dtype,
                                      device=device)

    def forward(self, x):
        return self.down_proj(fused.swiglu(self.gate_up_proj(x)))


class LlamaDecoderLayer(nn.Module):

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.input_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self.self_attn = LlamaAttention(cfg, device)
        self.post_attention_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self.mlp = LlamaMLP(cfg, device)

    def forward(self, h, cos, sin, position_ids=None, sp_attention=None):
        a = self.self_attn(self.input_layernorm(h), cos, sin, position_ids, sp_attention)
        x, h = self.post_attention_layernorm(a, residual=h)  # h <- h + a ; x = norm(h)
        return h + self.mlp(x)


class LlamaModel(nn.Module):

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.cfg = cfg
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=cfg.dtype, device=device)
        self.layers = nn.ModuleList([LlamaDecoderLayer(cfg, device) for _ in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self._rope_cache = {}
        self.sp_attention = None  # set by parallel.sp.enable_ulysses(model, group)

    def rope(self, seq_len, device):
        key = (seq_len, str(device))
        if key not in self._rope_cache:
            self._rope_cache = {key: fused.rope_tables(seq_len, self.cfg.head_dim, self.cfg.rope_theta, device,
                                                       self.cfg.rope_scaling)}
        return self._rope_cache[key]

    def forward(self, input_ids, position_ids=None):
        B, S = input_ids.shape
        max_pos = S if position_ids is None else None
        if max_pos is None:
            max_pos = self.cfg.max_position_embeddings if self.sp_attention is None else \
                max(self.cfg.max_position_embeddings, S * getattr(self.sp_attention, "world", 1))
        cos, sin = self.rope(max_pos, input_ids.device)
        h = self.embed_tokens(input_ids)
        # True: recompute every layer in backward; an int k: only the first k layers (partial checkpointing, to
        # spend exactly the HBM that is free); False/0: keep all activations
        gc = self.cfg.gradient_checkpointing
        n_ckpt = (len(self.layers) if gc is True else int(gc)) if (self.training and torch.is_grad_enabled()) else 0
        for i, layer in enumerate(self.layers):
            if i < n_ckpt:
                h = torch.utils.checkpoint.checkpoint(layer, h, cos, sin, position_ids, self.sp_attention,
                                                      use_reentrant=False)
            else:
                h = layer(h, cos, sin, position_ids, self.sp_attention)
        return self.norm(h)


class LMHead(nn.Module):
    """Output projection + (optionally) the fused chunked cross-entropy. Being a module, it is a ZeRO-3 unit:
    its weight is gathered by the unit hooks around ``forward`` (no "external parameter" access)."""

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size, dtype=cfg.dtype, device=device))
        self.chunk_tokens = cfg.loss_chunk_tokens

    def forward(self, h, targets=None, loss_tokens=None):
        if targets is None:
            return F.linear(h, self.weight)
        return fused.linear_cross_entropy(h, self.weight, targets, ignore_index=-100,
                                          chunk_tokens=self.chunk_tokens, reduction_tokens=loss_tokens)


class LlamaForCausalLM(nn.Module):
    """``forward(input_ids, labels)`` returns the mean next-token loss; without labels returns logits."""

    def __init__(self, cfg: LlamaConfig, device=None):
        super().__init__()
        self.cfg = cfg
        self.model = LlamaModel(cfg, device)
        self.lm_head = LMHead(cfg, device)
        if cfg.tie_word_embeddings:
            self.lm_head.weight = self.model.embed_tokens.weight
        self.reset_parameters()

    @torch.no_grad()
    def reset_parameters(self):
        std = self.cfg.init_std
        for name, p in self.named_parameters():
            if p.numel() == 0:
                continue
            if p.dim() >= 2:
                p.normal_(0.0, std)
            else:
                p.fill_(1.0)

    def forward(self, input_ids, labels=None, position_ids=None, shift_labels=True, loss_tokens=None):
        h = self.model(input_ids, position_ids)
        if labels is None:
            return self.lm_head(h)
        if shift_labels:
            # predict token t+1 from position t; the last position has no target
            tgt = torch.full_like(labels, -100)
            tgt[:, :-1] = labels[:, 1:]
        else:
            tgt = labels
        return self.lm_head(h, tgt, loss_tokens)


# ---- Hugging Face interop ---------------------------------------------------------------------------------
def hf_to_fused_state_dict(sd, cfg: LlamaConfig):
    """HF ``LlamaForCausalLM`` names -> fused (qkv_proj / gate_up_proj) names."""
    out = {}
    for k, v in sd.items():
        if any(s in k for s in (".q_proj.", ".k_proj.", ".v_proj.", ".gate_proj.", ".up_proj.")):
            continue
        out[k] = v
    for i in range(cfg.num_hidden_layers):
        p = f"model.layers.{i}."
        if p + "self_attn.q_proj.weight" in sd:
            out[p + "self_attn.qkv_proj.weight"] = torch.cat(
                [sd[p + "self_attn.q_proj.weight"], sd[p + "self_attn.k_proj.weight"], sd[p + "self_attn.v_proj.weight"]], 0)
            out[p + "mlp.gate_up_proj.weight"] = torch.cat([sd[p + "mlp.gate_proj.weight"], sd[p + "mlp.up_proj.weight"]], 0)
    return out


def fused_to_hf_state_dict(sd, cfg: LlamaConfig):
    out = {}
    q, kv = cfg.num_attention_heads * cfg.head_dim, cfg.num_key_value_heads * cfg.head_dim
    for k, v in sd.items():
        if k.endswith("self_attn.qkv_proj.weight"):
            base = k[:-len("qkv_proj.weight")]
            out[base + "q_proj.weight"], out[base + "k_proj.weight"], out[base + "v_proj.weight"] = v.split([q, kv, kv], 0)
        elif k.endswith("mlp.gate_up_proj.weight"):
            base = k[:-len("gate_up_proj.weight")]
            out[base + "gate_proj.weight"], out[base + "up_proj.weight"] = v.chunk(2, 0)
        else:
            out[k] = v
    return out
