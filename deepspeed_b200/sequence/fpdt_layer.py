"""FPDT / Ulysses-Offload: chunked attention with online-softmax merging and host offload of KV chunks.

Reference: ``sequence/fpdt_layer.py`` (``update_out_and_lse`` :41-77, ``_FPDTGPUAttentionImpl_`` :135-495,
offloading variant :545-1040, ``FPDT_Attention`` :1042, ``FPDT_FFN`` :1127, ``FPDT_LogitsLoss`` :1208,
``FPDT_InputConstruct`` :80-132). The local sequence is processed in chunks: every query chunk attends to all earlier
KV chunks (kept on the GPU, or in pinned host memory and double-buffered back on a side stream), partial results are
merged with the running log-sum-exp.
"""
import math

import torch
import torch.nn.functional as F

from .layer import heads_to_seq, seq_to_heads


def update_out_and_lse(out, lse, block_out, block_lse):
    """Merge two partial attention results (online softmax): out [B,S,H,D], lse [B,H,S]."""
    if out is None:
        return block_out.float(), block_lse.float()
    new_lse = torch.logaddexp(lse, block_lse)
    w_old = torch.exp(lse - new_lse).transpose(1, 2).unsqueeze(-1)
    w_new = torch.exp(block_lse - new_lse).transpose(1, 2).unsqueeze(-1)
    return out * w_old + block_out.float() * w_new, new_lse


def _attn_with_lse(q, k, v, causal):
    """q,k,v: [B,S,H,D] -> (out [B,S,H,D], lse [B,H,Sq]). On a GPU (bf16, head_dim 128, chunk a multiple of 128) this is
    the tcgen05 flash-attention forward/backward pair with the log-sum-exp as a differentiable output
    (``ops/attention.py``; reference: flash-attn ``_flash_attn_forward/_backward``, fpdt_layer.py:238-269); otherwise
    fp32 math that materialises the chunk's scores."""
    if q.is_cuda:
        from ..ops import attention as _A
        qc, kc, vc = (t if t.stride(-1) == 1 and t.stride(2) == t.shape[-1] else t.contiguous() for t in (q, k, v))
        if _A.lse_attention_supported(qc, kc, vc):
            return _A.flash_attention_with_lse(qc, kc, vc, causal=causal)
    qt, kt, vt = q.transpose(1, 2).float(), k.transpose(1, 2).float(), v.transpose(1, 2).float()
    if kt.shape[1] != qt.shape[1]:
        rep = qt.shape[1] // kt.shape[1]
        kt, vt = kt.repeat_interleave(rep, 1), vt.repeat_interleave(rep, 1)
    scores = qt @ kt.transpose(-1, -2) / math.sqrt(q.shape[-1])
    if causal:
        sq, sk = scores.shape[-2:]
        mask = torch.ones(sq, sk, dtype=torch.bool, device=q.device).tril(sk - sq)
        scores = scores.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(scores, dim=-1)
    out = torch.softmax(scores, dim=-1) @ vt
    return out.transpose(1, 2), lse


class SequenceChunk:
    """A KV chunk that may live in pinned host memory; ``get()`` brings it back (async on a side stream)."""

    def __init__(self, tensor, offload=False):
        self.device = tensor.device
        self.offloaded = bool(offload and tensor.is_cuda)
        if self.offloaded:
            self.host = torch.empty(tensor.shape, dtype=tensor.dtype, device="cpu", pin_memory=True)
            self.host.copy_(tensor, non_blocking=True)
            self.gpu = None
        else:
            self.gpu, self.host = tensor, None

    def get(self, stream=None):
        if not self.offloaded:
            return self.gpu
        if stream is not None:
            with torch.cuda.stream(stream):
                t = self.host.to(self.device, non_blocking=True)
            torch.cuda.current_stream().wait_stream(stream)
            return t
        return self.host.to(self.device, non_blocking=True)


def chunked_causal_attention(q, k, v, chunk_size, offload=False):
    """Blockwise causal attention over sequence chunks with online-softmax merge. q,k,v: [B,S,H,D]."""
    B, S, H, D = q.shape
    n = math.ceil(S / chunk_size)
    side = torch.cuda.Stream() if (offload and q.is_cuda) else None
    kv = [SequenceChunk(torch.stack([k[:, i * chunk_size:(i + 1) * chunk_size], v[:, i * chunk_size:(i + 1) * chunk_size]]),
                        offload) for i in range(n)]
    outs = []
    for i in range(n):
        qi = q[:, i * chunk_size:(i + 1) * chunk_size]
        out, lse = None, None
        for j in range(i + 1):
            kvj = kv[j].get(side)
            bo, bl = _attn_with_lse(qi, kvj[0], kvj[1], causal=(j == i))
            out, lse = update_out_and_lse(out, lse, bo, bl)
        outs.append(out.to(q.dtype))
    return torch.cat(outs, dim=1)


class FPDT_Attention(torch.nn.Module):
    """Sequence-parallel (Ulysses) + chunked attention. ``forward(q, k, v)`` with ``[B, S/P, H, D]`` inputs."""

    def __init__(self, sp_group=None, chunk_size=65536, enable_offload=False):
        super().__init__()
        self.sp_group, self.chunk_size, self.enable_offload = sp_group, chunk_size, enable_offload

    def forward(self, q, k, v):
        if self.sp_group is not None:
            q, k, v = seq_to_heads(q, self.sp_group), seq_to_heads(k, self.sp_group), seq_to_heads(v, self.sp_group)
        o = chunked_causal_attention(q, k, v, self.chunk_size, self.enable_offload)
        if self.sp_group is not None:
            o = heads_to_seq(o.contiguous(), self.sp_group)
        return o


class FPDT_FFN(torch.nn.Module):
    """Run an FFN over sequence chunks (activation memory of one chunk)."""

    def __init__(self, ffn, chunk_size=65536):
        super().__init__()
        self.ffn, self.chunk_size = ffn, chunk_size

    def forward(self, x):
        from .ulysses_sp import TiledMLP
        shards = max(1, math.ceil(x.shape[1] / self.chunk_size))
        return TiledMLP.apply(lambda m, t: m(t), self.ffn, x, shards, list(self.ffn.parameters()))


class FPDT_LogitsLoss(torch.nn.Module):
    """Chunked LM-head + loss (never materialises the full logits)."""

    def __init__(self, lm_head_weight, chunk_size=8192):
        super().__init__()
        self.weight, self.chunk_size = lm_head_weight, chunk_size

    def forward(self, hidden, labels):
        from ..ops.fused import linear_cross_entropy
        return linear_cross_entropy(hidden, self.weight, labels, chunk_tokens=self.chunk_size)


def FPDT_InputConstruct(tokens, labels, loss_mask, attention_mask, position_ids, sp_size, sp_rank, chunk_size=None):
    """Shard a batch on the sequence dimension for rank ``sp_rank`` (load-balanced interleave of chunk pairs so the
    causal work per rank is even: rank r takes chunks r and 2P-1-r)."""
    S = tokens.shape[1]
    if sp_size == 1:
        return tokens, labels, loss_mask, attention_mask, position_ids
    assert S % (2 * sp_size) == 0, "sequence length must be divisible by 2*sp_size for the balanced interleave"
    c = S // (2 * sp_size)
    idx = torch.cat([torch.arange(sp_rank * c, (sp_rank + 1) * c), torch.arange((2 * sp_size - 1 - sp_rank) * c,
                                                                              (2 * sp_size - sp_rank) * c)]).to(tokens.device)
    pick = lambda t: t.index_select(1, idx) if t is not None and torch.is_tensor(t) and t.dim() >= 2 else t  # noqa: E731
    if position_ids is None:
        position_ids = torch.arange(S, device=tokens.device).unsqueeze(0).expand_as(tokens)
    return pick(tokens), pick(labels), pick(loss_mask), attention_mask, pick(position_ids)


# ---- helper functions of the reference module (``sequence/fpdt_layer.py:33, 1115-1124``) ---------------------------
def apply_rotary_pos_emb_backward(grad_output, freqs_cos, freqs_sin):
    """Gradient of :func:`deepspeed_b200.sequence.layer.apply_rotary_pos_emb` w.r.t. its input: the adjoint of the
    half rotation is the rotation by the opposite angle."""
    rot = freqs_cos.shape[-1]
    g, rest = grad_output[..., :rot], grad_output[..., rot:]
    gs = g * freqs_sin
    g1, g2 = gs[..., :rot // 2], gs[..., rot // 2:]
    g = g * freqs_cos + torch.cat((g2, -g1), dim=-1)
    return g if rest.shape[-1] == 0 else torch.cat((g, rest), dim=-1)


def bias_gelu(x):
    """tanh-approximated GELU (the bias is added by the caller)."""
    return x * 0.5 * (1.0 + torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x)))


def bias_gelu_back(g, x):
    """``g * d/dx bias_gelu(x)``."""
    t = torch.tanh(0.79788456 * x * (1 + 0.044715 * x * x))
    return (0.5 * x * ((1 - t * t) * (0.79788456 + 0.1070322243 * x * x)) + 0.5 * (1 + t)) * g
