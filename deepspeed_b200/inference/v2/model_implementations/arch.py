"""One parameterised decoder that covers the FastGen model families (reference
``inference/v2/model_implementations/{llama_v2,mistral,mixtral,qwen,qwen_v2,qwen_v2_moe,phi,phi3,falcon,opt,...}``:
one container + policy + model class per family, ~600 LoC each). The families differ in a handful of switches
(normalisation, gated or plain MLP, biases, partial rotary / learned positions, parallel attention+MLP block, MoE
with or without a shared expert); :class:`ArchSpec` names those switches and :class:`GenericDecoder` implements them
once. Attention itself is delegated to ``attn_fn`` so that the same module serves dense evaluation and the ragged
engine with its blocked KV cache."""
from dataclasses import dataclass, field
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ....ops import fastgen

from ....ops import fused


@dataclass
class ArchSpec:
    family: str = "llama"
    vocab_size: int = 32000
    hidden_size: int = 4096
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 32
    head_dim: Optional[int] = None
    intermediate_size: int = 11008
    norm: str = "rms"                 # rms | ln
    norm_eps: float = 1e-5
    act: str = "silu"                 # silu | gelu | gelu_tanh | relu
    gated_mlp: bool = True
    qkv_bias: bool = False
    o_bias: bool = False
    mlp_bias: bool = False
    pos: str = "rope"                 # rope | learned
    rotary_dim: Optional[int] = None  # partial rotary (phi); None = whole head
    rope_theta: float = 10000.0
    rope_scaling: Optional[dict] = None
    pos_offset: int = 0               # OPT stores learned positions shifted by 2
    max_position_embeddings: int = 4096
    parallel_block: bool = False      # falcon / phi: attention and MLP read the same normed input
    parallel_two_norms: bool = False  # falcon new decoder architecture: ln_attn + ln_mlp
    final_norm: bool = True
    qk_norm: bool = False             # exaone4: RMSNorm over head_dim of q and k before RoPE
    sublayer_post_norm: bool = False  # exaone4: h += norm(attn(h)); h += norm(mlp(h)), no pre-norm
    lm_head_bias: bool = False
    tie_word_embeddings: bool = False
    num_experts: int = 0
    num_experts_per_tok: int = 2
    norm_topk_prob: bool = True
    moe_intermediate_size: Optional[int] = None
    shared_expert_intermediate_size: int = 0  # qwen2-moe: always-on expert with a sigmoid gate
    dtype: torch.dtype = torch.bfloat16
    extra: dict = field(default_factory=dict)

    def __post_init__(self):
        if self.head_dim is None:
            self.head_dim = self.hidden_size // self.num_attention_heads


def _act(name):
    return {"silu": F.silu, "gelu": F.gelu, "relu": F.relu,
            "gelu_tanh": lambda x: F.gelu(x, approximate="tanh")}[name]


class _Norm(nn.Module):

    def __init__(self, spec, device=None):
        super().__init__()
        self.kind, self.eps = spec.norm, spec.norm_eps
        self.weight = nn.Parameter(torch.ones(spec.hidden_size, dtype=spec.dtype, device=device))
        self.bias = nn.Parameter(torch.zeros(spec.hidden_size, dtype=spec.dtype, device=device)) \
            if spec.norm == "ln" else None

    def forward(self, x):
        if self.kind == "rms":
            return fused.rms_norm(x, self.weight, self.eps)
        return fused.layer_norm(x, self.weight, self.bias, self.eps)


class _MLP(nn.Module):

    def __init__(self, spec, inter, device=None):
        super().__init__()
        h, kw = spec.hidden_size, dict(dtype=spec.dtype, device=device)
        self.gated, self.act_name = spec.gated_mlp, spec.act
        self.up = nn.Linear(h, inter * (2 if spec.gated_mlp else 1), bias=spec.mlp_bias, **kw)  # [gate | up] fused
        self.down = nn.Linear(inter, h, bias=spec.mlp_bias, **kw)

    def forward(self, x):
        if self.gated and self.act_name == "silu" and self.up.bias is None:
            return self.down(fused.swiglu(self.up(x)))
        # bias add and (gated) activation in one pass over the projection's output (csrc/fastgen_ops.cu)
        u = F.linear(x, self.up.weight)
        if self.gated:
            return self.down(fastgen.gated_act(u, self.up.bias, self.act_name))
        return self.down(fastgen.bias_act(u, self.up.bias, self.act_name))


class _MoE(nn.Module):
    """Token-choice top-k router over dense experts (inference: experts are evaluated on their routed tokens only)."""

    def __init__(self, spec, device=None):
        super().__init__()
        kw = dict(dtype=spec.dtype, device=device)
        inter = spec.moe_intermediate_size or spec.intermediate_size
        self.top_k, self.norm_topk = spec.num_experts_per_tok, spec.norm_topk_prob
        self.gate = nn.Linear(spec.hidden_size, spec.num_experts, bias=False, **kw)
        self.experts = nn.ModuleList([_MLP(spec, inter, device) for _ in range(spec.num_experts)])
        self.shared = None
        if spec.shared_expert_intermediate_size:
            self.shared = _MLP(spec, spec.shared_expert_intermediate_size, device)
            self.shared_gate = nn.Linear(spec.hidden_size, 1, bias=False, **kw)

    def _stack_experts(self):
        """Expert weights side by side for the grouped GEMM: ``w13 [E, 2I, H]`` / ``w2 [E, H, I]``. The per-expert
        ``nn.Linear`` weights become views of the stacks, so nothing is held twice."""
        w13 = torch.stack([e.up.weight.data for e in self.experts]).contiguous()
        w2 = torch.stack([e.down.weight.data for e in self.experts]).contiguous()
        for i, e in enumerate(self.experts):
            e.up.weight.data, e.down.weight.data = w13[i], w2[i]
        self._w13, self._w2 = w13, w2

    def _grouped_ok(self, x):
        import os
        e0 = self.experts[0]
        return (x.is_cuda and os.environ.get("DSB200_FASTGEN_GROUPED_MOE", "1") != "0" and e0.gated
                and e0.act_name == "silu" and e0.up.bias is None and e0.down.bias is None
                and x.dtype in (torch.bfloat16, torch.float16) and x.shape[-1] % 8 == 0 and fastgen._native(x))

    def _forward_grouped(self, x):
        """Gating kernel -> scatter to expert-contiguous rows -> two grouped GEMMs around SwiGLU -> weighted gather.
        One host read (the expert counts) per layer instead of one ``nonzero`` per expert."""
        from ....ops.grouped_gemm import grouped_mm
        if getattr(self, "_w13", None) is None or self._w13.device != x.device:
            self._stack_experts()
        scores, experts, counts, slot = fastgen.topk_gating(self.gate(x), self.top_k, self.norm_topk, with_slots=True)
        starts = torch.cumsum(counts, 0, dtype=torch.int32) - counts
        rows = starts[experts.long()] + slot
        n_rows = x.shape[0] * self.top_k
        xs = fastgen.moe_scatter(x, rows, n_rows)
        cl = counts.tolist()
        h = fused.swiglu(grouped_mm(xs, self._w13, cl))
        return fastgen.moe_gather(grouped_mm(h, self._w2, cl), rows, scores)

    def forward(self, x):
        shape = x.shape
        x = x.reshape(-1, shape[-1])
        if self._grouped_ok(x):
            out = self._forward_grouped(x.contiguous())
            if self.shared is not None:
                out = out + torch.sigmoid(self.shared_gate(x)) * self.shared(x)
            return out.view(shape)
        probs = torch.softmax(self.gate(x).float(), dim=-1)
        w, idx = probs.topk(self.top_k, dim=-1)
        if self.norm_topk:
            w = w / w.sum(-1, keepdim=True)
        out = torch.zeros_like(x)
        for e, expert in enumerate(self.experts):
            tok, slot = (idx == e).nonzero(as_tuple=True)
            if tok.numel():
                out.index_add_(0, tok, expert(x[tok]) * w[tok, slot].unsqueeze(-1).to(x.dtype))
        if self.shared is not None:
            out = out + torch.sigmoid(self.shared_gate(x)) * self.shared(x)
        return out.view(shape)


class _Attention(nn.Module):

    def __init__(self, spec, device=None):
        super().__init__()
        kw = dict(dtype=spec.dtype, device=device)
        self.nh, self.nkv, self.hd = spec.num_attention_heads, spec.num_key_value_heads, spec.head_dim
        self.qkv_proj = nn.Linear(spec.hidden_size, (self.nh + 2 * self.nkv) * self.hd, bias=spec.qkv_bias, **kw)
        self.o_proj = nn.Linear(self.nh * self.hd, spec.hidden_size, bias=spec.o_bias, **kw)
        if spec.qk_norm:
            self.q_norm = nn.Parameter(torch.ones(self.hd, **kw))
            self.k_norm = nn.Parameter(torch.ones(self.hd, **kw))


class _Layer(nn.Module):

    def __init__(self, spec, device=None):
        super().__init__()
        self.spec = spec
        if not spec.sublayer_post_norm:
            self.input_layernorm = _Norm(spec, device)
        self.self_attn = _Attention(spec, device)
        if not spec.parallel_block or spec.parallel_two_norms:
            self.post_attention_layernorm = _Norm(spec, device)
        if spec.sublayer_post_norm:
            self.post_feedforward_layernorm = _Norm(spec, device)
        self.mlp = _MoE(spec, device) if spec.num_experts else _MLP(spec, spec.intermediate_size, device)


def _rope_partial(x, cos, sin, pos, rotary_dim):
    """x: [T, heads, head_dim]; rotate the first ``rotary_dim`` channels (half-rotation convention)."""
    rot, rest = x[..., :rotary_dim], x[..., rotary_dim:]
    c, s = cos[pos].unsqueeze(1).to(x.dtype), sin[pos].unsqueeze(1).to(x.dtype)   # [T, 1, rotary_dim/2]
    x1, x2 = rot[..., :rotary_dim // 2], rot[..., rotary_dim // 2:]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s, rest], dim=-1)


class GenericDecoder(nn.Module):
    """``forward_tokens(ids [T], pos [T], attn_fn)`` -> hidden states; ``attn_fn(layer, q, k, v)`` receives
    [T, heads, head_dim] tensors (positions already encoded) and returns [T, heads * head_dim]."""

    def __init__(self, spec: ArchSpec, device=None):
        super().__init__()
        self.spec = self.cfg = spec
        kw = dict(dtype=spec.dtype, device=device)
        self.embed_tokens = nn.Embedding(spec.vocab_size, spec.hidden_size, **kw)
        if spec.pos == "learned":
            self.embed_positions = nn.Embedding(spec.max_position_embeddings + spec.pos_offset, spec.hidden_size, **kw)
        self.layers = nn.ModuleList([_Layer(spec, device) for _ in range(spec.num_hidden_layers)])
        self.norm = _Norm(spec, device) if spec.final_norm else None
        self.lm_head = nn.Linear(spec.hidden_size, spec.vocab_size, bias=spec.lm_head_bias, **kw)
        if spec.tie_word_embeddings:
            self.lm_head.weight = self.embed_tokens.weight
        self._rope = None

    def rope(self, device):
        if self._rope is None or self._rope[0].device != device:
            rd = self.spec.rotary_dim or self.spec.head_dim
            self._rope = fused.rope_tables(self.spec.max_position_embeddings, rd, self.spec.rope_theta, device,
                                           self.spec.rope_scaling)
        return self._rope

    def forward_tokens(self, ids, pos, attn_fn):
        spec = self.spec
        if spec.pos == "learned":  # token + position lookup in one kernel (csrc/fastgen_ops.cu embed_rows)
            h = fastgen.embed(ids, self.embed_tokens.weight, pos, self.embed_positions.weight, spec.pos_offset)
        else:
            h = fastgen.embed(ids, self.embed_tokens.weight)
            cos, sin = self.rope(ids.device)
            rd = spec.rotary_dim or spec.head_dim
        for li, layer in enumerate(self.layers):
            a = layer.self_attn
            x = h if spec.sublayer_post_norm else layer.input_layernorm(h)
            q, k, v = a.qkv_proj(x).split([a.nh * a.hd, a.nkv * a.hd, a.nkv * a.hd], dim=-1)
            q, k, v = q.reshape(-1, a.nh, a.hd), k.reshape(-1, a.nkv, a.hd), v.reshape(-1, a.nkv, a.hd)
            if spec.qk_norm:
                q, k = fused.rms_norm(q, a.q_norm, spec.norm_eps), fused.rms_norm(k, a.k_norm, spec.norm_eps)
            if spec.pos == "rope":
                q, k = _rope_partial(q, cos, sin, pos, rd), _rope_partial(k, cos, sin, pos, rd)
            attn_out = a.o_proj(attn_fn(li, q, k, v))
            if spec.sublayer_post_norm:
                h = h + layer.post_attention_layernorm(attn_out)
                h = h + layer.post_feedforward_layernorm(layer.mlp(h))
            elif spec.parallel_block:
                m_in = layer.post_attention_layernorm(h) if spec.parallel_two_norms else x
                h = h + attn_out + layer.mlp(m_in)
            else:
                h = h + attn_out
                h = h + layer.mlp(layer.post_attention_layernorm(h))
        return self.norm(h) if self.norm is not None else h

    def forward(self, input_ids):
        """Dense causal evaluation of [B, S] token ids -> logits [B, S, vocab] (tests / prompt scoring)."""
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device).repeat(B)

        def attn_fn(li, q, k, v):
            nh, nkv, hd = q.shape[1], k.shape[1], q.shape[2]
            q = q.view(B, S, nh, hd).transpose(1, 2)
            k = k.view(B, S, nkv, hd).transpose(1, 2).repeat_interleave(nh // nkv, 1)
            v = v.view(B, S, nkv, hd).transpose(1, 2).repeat_interleave(nh // nkv, 1)
            o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
            return o.transpose(1, 2).reshape(B * S, nh * hd)

        h = self.forward_tokens(input_ids.reshape(-1), pos, attn_fn)
        return self.lm_head(h).view(B, S, -1)
