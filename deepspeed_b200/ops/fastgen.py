"""FastGen / inference row operations on the native kernels of ``csrc/fastgen_ops.cu``.

Reference: ``inference/v2/kernels/ragged_ops/{embed,logits_gather,moe_scatter,moe_gather,top_k_gating}`` and
``core_ops/{bias_activations,gated_activations}`` (v1: ``fused_bias_gelu`` / ``fused_bias_relu`` /
``fused_gate_activation``). Every function has the same meaning on a CPU tensor (plain PyTorch) - the numerics tests
compare the two."""
import ctypes

import torch
import torch.nn.functional as F

from . import kernels as K

ACTS = {"identity": 0, "relu": 1, "gelu": 2, "gelu_tanh": 3, "silu": 4}
_declared = False


def _lib():
    global _declared
    L = K.lib()
    if not _declared:
        P, I, L64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
        L.dsb200_embed_rows.argtypes = [P, P, P, P, I, P, L64, I, I, P]
        L.dsb200_bias_act.argtypes = [P, P, P, L64, I, I, I, P]
        L.dsb200_gated_act.argtypes = [P, P, P, L64, I, I, I, P]
        L.dsb200_topk_gating.argtypes = [P, I, I, I, I, P, P, P, P, P]
        L.dsb200_moe_scatter.argtypes = [P, P, I, I, P, I, I, P]
        L.dsb200_moe_gather.argtypes = [P, P, P, I, I, P, I, I, P]
        _declared = True
    return L


def _native(x, cols=None):
    cols = x.shape[-1] if cols is None else cols
    return (x.is_cuda and K.available() and hasattr(K.lib(), "dsb200_embed_rows")
            and x.dtype in (torch.bfloat16, torch.float16) and cols % 8 == 0 and x.data_ptr() % 16 == 0)


def _torch_act(name):
    return {"identity": lambda t: t, "relu": F.relu, "gelu": F.gelu, "silu": F.silu,
            "gelu_tanh": lambda t: F.gelu(t, approximate="tanh")}[name]


def embed(ids, table, positions=None, pos_table=None, pos_offset=0):
    """``table[ids] (+ pos_table[positions + pos_offset])`` for a flat (ragged) batch of token ids."""
    flat = ids.reshape(-1)
    if _native(table) and table.is_contiguous() and (pos_table is None or (_native(pos_table)
                                                                             and pos_table.is_contiguous())):
        out = torch.empty(flat.numel(), table.shape[1], dtype=table.dtype, device=table.device)
        idx = flat.to(torch.int64).contiguous()
        pos = positions.reshape(-1).to(torch.int64).contiguous() if pos_table is not None else None
        K._check(_lib().dsb200_embed_rows(K._ptr(table), K._ptr(idx), K._ptr(pos_table), K._ptr(pos), int(pos_offset),
                                          K._ptr(out), flat.numel(), table.shape[1], K._dt(table), K._stream()),
                 "embed_rows")
        return out.view(*ids.shape, table.shape[1])
    out = F.embedding(ids, table)
    if pos_table is not None:
        out = out + F.embedding(positions + pos_offset, pos_table)
    return out


def logits_gather(hidden, last_index):
    """Rows ``hidden[last_index]``: the last token of every sequence of a ragged batch, input of the LM head."""
    return embed(last_index, hidden) if hidden.dim() == 2 else hidden.reshape(-1, hidden.shape[-1])[last_index]


def bias_act(x, bias=None, act="gelu"):
    """``act(x + bias)`` in one pass."""
    if _native(x) and x.is_contiguous() and (bias is None or (bias.dtype == x.dtype and bias.data_ptr() % 16 == 0)):
        y = torch.empty_like(x)
        K._check(_lib().dsb200_bias_act(K._ptr(x), K._ptr(bias), K._ptr(y), x.numel() // x.shape[-1], x.shape[-1],
                                        ACTS[act], K._dt(x), K._stream()), "bias_act")
        return y
    return _torch_act(act)(x if bias is None else x + bias)


def gated_act(x, bias=None, act="silu"):
    """``act(gate + b_gate) * (up + b_up)`` for ``x = [gate | up]`` along the last dim (SwiGLU, GeGLU, ReGLU)."""
    inter = x.shape[-1] // 2
    if _native(x, inter) and x.is_contiguous() and (bias is None or (bias.dtype == x.dtype
                                                                       and bias.data_ptr() % 16 == 0)):
        y = torch.empty(*x.shape[:-1], inter, dtype=x.dtype, device=x.device)
        K._check(_lib().dsb200_gated_act(K._ptr(x), K._ptr(bias), K._ptr(y), x.numel() // x.shape[-1], inter, ACTS[act],
                                         K._dt(x), K._stream()), "gated_act")
        return y
    if bias is not None:
        x = x + bias
    g, u = x.chunk(2, dim=-1)
    return _torch_act(act)(g) * u


def topk_gating(logits, k, renormalize=True, with_slots=False):
    """Softmax over the experts, the ``k`` best per token. Returns ``(scores [T, k] fp32, experts [T, k] int32,
    counts [E] int32)`` and, with ``with_slots``, the arrival order of every assignment inside its expert."""
    T, E = logits.shape
    if logits.is_cuda and K.available() and hasattr(K.lib(), "dsb200_topk_gating") and E <= 1024 and k <= 16:
        lg = logits.float().contiguous()
        scores = torch.empty(T, k, dtype=torch.float32, device=logits.device)
        experts = torch.empty(T, k, dtype=torch.int32, device=logits.device)
        counts = torch.zeros(E, dtype=torch.int32, device=logits.device)
        slot = torch.empty(T, k, dtype=torch.int32, device=logits.device) if with_slots else None
        K._check(_lib().dsb200_topk_gating(K._ptr(lg), T, E, k, int(bool(renormalize)), K._ptr(scores), K._ptr(experts),
                                           K._ptr(counts), K._ptr(slot), K._stream()), "topk_gating")
        return (scores, experts, counts, slot) if with_slots else (scores, experts, counts)
    probs = torch.softmax(logits.float(), dim=-1)
    scores, experts = probs.topk(k, dim=-1)
    if renormalize:
        scores = scores / scores.sum(-1, keepdim=True).clamp(min=1e-20)
    counts = torch.bincount(experts.reshape(-1), minlength=E).to(torch.int32)
    experts = experts.to(torch.int32)
    if not with_slots:
        return scores, experts, counts
    flat = experts.reshape(-1).long()
    order = torch.argsort(flat, stable=True)
    ranks = torch.empty_like(flat)
    starts = torch.cumsum(counts.long(), 0) - counts.long()
    ranks[order] = torch.arange(flat.numel(), device=flat.device) - starts[flat[order]]
    return scores, experts, counts, ranks.view(T, k).to(torch.int32)


def moe_scatter(x, dst_rows, num_rows):
    """``out[dst_rows[t, j]] = x[t]``: tokens copied to their expert-contiguous rows (``dst < 0``: dropped)."""
    T, k = dst_rows.shape
    out = torch.zeros(num_rows, x.shape[1], dtype=x.dtype, device=x.device)
    if _native(x) and x.is_contiguous():
        d = dst_rows.to(torch.int32).contiguous()
        K._check(_lib().dsb200_moe_scatter(K._ptr(x), K._ptr(d), T, k, K._ptr(out), x.shape[1], K._dt(x), K._stream()),
                 "moe_scatter")
        return out
    flat = dst_rows.reshape(-1).long()
    keep = flat >= 0
    tok = torch.arange(T, device=x.device).repeat_interleave(k)
    out[flat[keep]] = x[tok[keep]]
    return out


def moe_gather(expert_out, src_rows, weights):
    """``out[t] = sum_j weights[t, j] * expert_out[src_rows[t, j]]`` (``src < 0``: dropped)."""
    T, k = src_rows.shape
    if _native(expert_out) and expert_out.is_contiguous():
        out = torch.empty(T, expert_out.shape[1], dtype=expert_out.dtype, device=expert_out.device)
        s = src_rows.to(torch.int32).contiguous()
        w = weights.float().contiguous()
        K._check(_lib().dsb200_moe_gather(K._ptr(expert_out), K._ptr(s), K._ptr(w), T, k, K._ptr(out),
                                          expert_out.shape[1], K._dt(expert_out), K._stream()), "moe_gather")
        return out
    idx = src_rows.long().clamp(min=0)
    picked = expert_out[idx.reshape(-1)].view(T, k, -1).float()
    w = (weights.float() * (src_rows >= 0)).unsqueeze(-1)
    return (picked * w).sum(1).to(expert_out.dtype)
