"""Native flash-attention front-end: the hand-written tcgen05 kernels of ``csrc/attention_fwd.cu`` /
``csrc/attention_bwd.cu`` behind a ``torch.autograd.Function``.

Layout is the projection layout ``[batch, seq, heads, 128]`` (views of a fused QKV projection are consumed in
place through strided TMA tensor maps). ``flash_attention_forward`` also returns the log-sum-exp, which the
chunked / sequence-parallel callers (FPDT ``sequence/fpdt_layer.py`` online-softmax merging) need and the framework
SDPA call does not expose. Reference call sites: ``sequence/fpdt_layer.py:238-269`` (flash-attn forward/backward with
LSE), ``runtime/sequence_parallel/ulysses_sp.py:50`` (HF attention functions).
"""
import ctypes
import math
import os

import torch

from . import kernels as K

_DECLARED = False


def _lib():
    global _DECLARED
    L = K.lib()
    if not _DECLARED:
        P, I, L64, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
        L.dsb200_flash_attn_fwd.argtypes = [P, P, P, P, P, I, I, I, I, I, L64, L64, L64, L64, L64, L64, L64, L64, F, I, P]
        L.dsb200_flash_attn_fwd.restype = I
        L.dsb200_flash_attn_fwd_paged.argtypes = [P, P, P, P, P, P, P, P, I, I, I, I, I, I, I, L64, L64, F, P]
        L.dsb200_flash_attn_fwd_paged.restype = I
        L.dsb200_flash_decode_paged.argtypes = [P, P, P, P, P, P, P, I, I, I, I, I, I, F, P]
        L.dsb200_flash_decode_paged.restype = I
        if hasattr(L, "dsb200_flash_attn_fwd_bias"):
            L.dsb200_flash_attn_fwd_bias.argtypes = [P] * 7 + [I, I, I, I] + [L64] * 12 + [I, F, P]
            L.dsb200_flash_attn_fwd_bias.restype = I
        if hasattr(L, "dsb200_flash_attn_bwd"):
            L.dsb200_flash_attn_bwd.argtypes = [P] * 11 + [I, I, I, I] + [L64] * 16 + [F, I, P]
            L.dsb200_flash_attn_bwd.restype = I
        _DECLARED = True
    return L


def _ok_layout(t):
    """[B, S, H, 128] bf16 with the head dim contiguous and heads packed 128 apart (strided rows / batches allowed)."""
    return (t.dim() == 4 and t.dtype == torch.bfloat16 and t.shape[-1] == 128 and t.stride(3) == 1
            and (t.shape[2] == 1 or t.stride(2) == 128) and t.stride(1) % 8 == 0 and t.stride(0) % 8 == 0
            and t.data_ptr() % 16 == 0)


def supported(q, k, v):
    """True when the hand-written sm_100a attention kernel can serve these operands."""
    if not (q.is_cuda and K.available() and hasattr(K.lib(), "dsb200_flash_attn_fwd")):
        return False
    return (_ok_layout(q) and _ok_layout(k) and _ok_layout(v) and k.shape == v.shape and q.shape[0] == k.shape[0]
            and q.shape[2] % k.shape[2] == 0)


def enabled_by_default():
    """Training default: DSB200_NATIVE_ATTN=1 forces the native kernels, =0 the library SDPA."""
    return os.environ.get("DSB200_NATIVE_ATTN", "0") == "1"


def flash_attention_forward(q, k, v, causal=True, scale=None, out=None, want_lse=True):
    """q [B,Sq,H,128], k/v [B,Skv,Hkv,128] -> (o [B,Sq,H,128] bf16, lse [B,H,Sq] fp32 or None)."""
    assert supported(q, k, v), "flash_attention_forward: unsupported operands (bf16, head_dim 128, packed heads)"
    B, Sq, H, D = q.shape
    Skv, Hkv = k.shape[1], k.shape[2]
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if out is None:
        out = torch.empty((B, Sq, H, D), dtype=torch.bfloat16, device=q.device)
    lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device) if want_lse else None
    rc = _lib().dsb200_flash_attn_fwd(
        K._ptr(q), K._ptr(k), K._ptr(v), K._ptr(out), K._ptr(lse), B, Sq, Skv, H, Hkv, q.stride(1), q.stride(0),
        k.stride(1), k.stride(0), v.stride(1), v.stride(0), out.stride(1), out.stride(0), float(scale),
        1 if causal else 0, K._stream())
    K._check(rc, "flash_attn_fwd")
    return out, lse


def bias_attention_supported(q):
    return q.is_cuda and K.available() and hasattr(K.lib(), "dsb200_flash_attn_fwd_bias")


def flash_attention_bias_forward(q, k, v, bias1=None, bias2=None, bias2_div=1, scale=None):
    """Non-causal attention with additive biases on the tcgen05 kernel (``csrc/attention_fwd.cu``, BIAS variant).

    q, k, v: [B, S, H, D] with D <= 128 (bf16 / fp16 / fp32; computed in bf16, heads zero-padded to 128);
    ``bias1`` [B, Skv] added to every row and head; ``bias2`` [B / bias2_div, H, Sq, Skv] shared by ``bias2_div``
    consecutive batch entries. Returns (o [B, S, H, D] in q's dtype, lse [B, H, Sq] fp32, natural log)."""
    B, Sq, H, D = q.shape
    Skv = k.shape[1]
    assert D <= 128 and k.shape == v.shape and k.shape[2] == H
    if scale is None:
        scale = 1.0 / math.sqrt(D)

    def pad(t):
        t = t.to(torch.bfloat16)
        if D == 128 and t.is_contiguous():
            return t
        out = torch.zeros(t.shape[:-1] + (128, ), dtype=torch.bfloat16, device=t.device)
        out[..., :D] = t
        return out

    def bias_rows(b, last):
        """bf16, last dim padded to a multiple of 8 so that every row is 16-byte aligned."""
        if b is None:
            return None
        b = b.to(torch.bfloat16)
        padn = (-last) % 8
        if padn:
            b = torch.nn.functional.pad(b, (0, padn))
        return b.contiguous()

    qp, kp, vp = pad(q), pad(k), pad(v)
    out = torch.empty((B, Sq, H, 128), dtype=torch.bfloat16, device=q.device)
    lse = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device)
    b1 = bias_rows(bias1.reshape(B, Skv), Skv) if bias1 is not None else None
    b2 = bias_rows(bias2.reshape(-1, H, Sq, Skv), Skv) if bias2 is not None else None
    if b2 is not None:
        assert b2.shape[0] * bias2_div == B, "bias2 batch x bias2_div must equal the attention batch"
    rc = _lib().dsb200_flash_attn_fwd_bias(
        K._ptr(qp), K._ptr(kp), K._ptr(vp), K._ptr(out), K._ptr(lse), K._ptr(b1), K._ptr(b2), B, Sq, Skv, H,
        qp.stride(1), qp.stride(0), kp.stride(1), kp.stride(0), vp.stride(1), vp.stride(0), out.stride(1),
        out.stride(0), b1.stride(0) if b1 is not None else 0, b2.stride(0) if b2 is not None else 0,
        b2.stride(1) if b2 is not None else 0, b2.stride(2) if b2 is not None else 0, int(bias2_div), float(scale),
        K._stream())
    K._check(rc, "flash_attn_fwd_bias")
    return out[..., :D].to(q.dtype), lse


def paged_prefill_supported(q, kv_layer):
    """q [T, H, 128] bf16 (heads packed), kv_layer [num_blocks, 2, 128, Hkv, 128] contiguous bf16."""
    return (q.is_cuda and K.available() and hasattr(K.lib(), "dsb200_flash_attn_fwd_paged") and q.dtype == torch.bfloat16
            and q.dim() == 3 and q.shape[-1] == 128 and q.stride(2) == 1 and q.stride(1) == 128 and q.stride(0) % 8 == 0
            and q.data_ptr() % 16 == 0 and kv_layer.dim() == 5 and kv_layer.shape[2] == 128 and kv_layer.shape[4] == 128
            and kv_layer.is_contiguous() and q.shape[1] % kv_layer.shape[3] == 0)


def flash_attention_paged_prefill(q, kv_layer, cu_q, q_len, kv_len, block_table, max_q_len, scale=None, out=None):
    """Causal attention of a ragged prefill batch over a blocked KV cache on the tcgen05 kernel.

    ``q`` [T_total, H, 128]: sequences packed back to back, sequence s = rows [cu_q[s], cu_q[s] + q_len[s]); its keys
    and values (including this chunk's, already appended) are the first kv_len[s] positions of blocks
    ``block_table[s]`` of ``kv_layer`` [num_blocks, 2, 128, Hkv, 128]. ``cu_q`` / ``q_len`` / ``kv_len`` /
    ``block_table`` are int32 device tensors. Rows of ``out`` outside the listed sequences are left untouched."""
    assert paged_prefill_supported(q, kv_layer)
    T, H, D = q.shape
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if out is None:
        out = torch.empty((T, H, D), dtype=torch.bfloat16, device=q.device)
    rc = _lib().dsb200_flash_attn_fwd_paged(
        K._ptr(q), K._ptr(kv_layer), K._ptr(out), None, K._ptr(cu_q), K._ptr(q_len), K._ptr(kv_len), K._ptr(block_table),
        cu_q.numel(), T, int(max_q_len), H, kv_layer.shape[3], kv_layer.shape[0], block_table.shape[1], q.stride(0),
        out.stride(0), float(scale), K._stream())
    K._check(rc, "flash_attn_fwd_paged")
    return out


_DECODE_SCRATCH = {}


def flash_decode_paged(q, kv_layer, kv_len, block_table, max_context, scale=None, out=None, splits=None):
    """Decode attention (one query token per sequence) over a blocked KV cache on the tcgen05 kernel: the query heads
    that share a KV head form the rows of the MMA tile, K / V blocks arrive by TMA through the block table and the key
    range is split over CTAs (flash-decoding) so that small batches still cover all SMs.

    ``q`` [T, Hq, 128] contiguous bf16; ``kv_layer`` [num_blocks, 2, 128, Hkv, 128]; ``kv_len`` [T] int32 (keys visible
    to each token, its own included); ``block_table`` [T, max_blocks] int32. Returns ``out`` [T, Hq, 128]."""
    assert paged_prefill_supported(q, kv_layer) and q.is_contiguous()
    T, Hq, D = q.shape
    Hkv = kv_layer.shape[3]
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if out is None:
        out = torch.empty((T, Hq, D), dtype=torch.bfloat16, device=q.device)
    n_blocks_max = max(1, (int(max_context) + 127) // 128)
    if splits is None:
        splits = max(1, min(n_blocks_max, -(-2 * 148 // max(1, T * Hkv)), 64))
    po = pl = None
    if splits > 1:
        need = T * Hq * splits * (D + 1)
        buf = _DECODE_SCRATCH.get(q.device)
        if buf is None or buf.numel() < need:
            buf = _DECODE_SCRATCH[q.device] = torch.empty(need, dtype=torch.float32, device=q.device)
        po, pl = buf[:T * Hq * splits * D], buf[T * Hq * splits * D:need]
    rc = _lib().dsb200_flash_decode_paged(K._ptr(q), K._ptr(kv_layer), K._ptr(out), K._ptr(kv_len), K._ptr(block_table),
                                          K._ptr(po), K._ptr(pl), T, Hq, Hkv, kv_layer.shape[0], block_table.shape[1],
                                          int(splits), float(scale), K._stream())
    K._check(rc, "flash_decode_paged")
    return out


def _reference_backward(q, k, v, do, causal, scale):
    """Library fallback for the backward pass (recomputes through the framework SDPA)."""
    with torch.enable_grad():
        qd, kd, vd = (t.detach().requires_grad_(True) for t in (q, k, v))
        o = torch.nn.functional.scaled_dot_product_attention(qd.transpose(1, 2), kd.transpose(1, 2), vd.transpose(1, 2),
                                                             is_causal=causal, scale=scale,
                                                             enable_gqa=q.shape[2] != k.shape[2])
        o.backward(do.transpose(1, 2))
    return qd.grad, kd.grad, vd.grad


def _native_backward_available():
    return K.available() and hasattr(K.lib(), "dsb200_flash_attn_bwd") and os.environ.get("DSB200_NATIVE_ATTN_BWD", "1") == "1"


def native_backward_supported(q, k, v, o, do, lse):
    return (_native_backward_available() and q.shape[1] == k.shape[1] and q.shape[1] % 128 == 0 and lse is not None
            and _ok_layout(o) and _ok_layout(do) and supported(q, k, v))


def flash_attention_backward(q, k, v, o, lse, do, causal=True, scale=None, dlse=None):
    """Gradients (dq, dk, dv), contiguous [B,S,H,128] / [B,S,Hkv,128], on the tcgen05 backward kernels
    (``csrc/attention_bwd.cu``) for Sq == Skv a multiple of 128; otherwise through the library SDPA. ``dlse``: gradient
    w.r.t. the log-sum-exp output [B,H,S] (chunked attention merges partial results through it) or None."""
    B, Sq, H, D = q.shape
    Skv, Hkv = k.shape[1], k.shape[2]
    if scale is None:
        scale = 1.0 / math.sqrt(D)
    if not native_backward_supported(q, k, v, o, do, lse):
        if dlse is not None:
            raise RuntimeError("flash_attention_backward: a gradient w.r.t. the LSE output needs the native kernels "
                               "(bf16, head_dim 128, Sq == Skv multiple of 128)")
        return _reference_backward(q, k, v, do, causal, scale)
    L = _lib()
    dq = torch.empty((B, Sq, H, D), dtype=torch.bfloat16, device=q.device)
    dk = torch.empty((B, Skv, Hkv, D), dtype=torch.bfloat16, device=q.device)
    dv = torch.empty((B, Skv, Hkv, D), dtype=torch.bfloat16, device=q.device)
    delta = torch.empty((B, H, Sq), dtype=torch.float32, device=q.device)
    rc = L.dsb200_flash_attn_bwd(
        K._ptr(q), K._ptr(k), K._ptr(v), K._ptr(o), K._ptr(do), K._ptr(lse),
        K._ptr(dlse.contiguous().float() if dlse is not None else None), K._ptr(delta), K._ptr(dq), K._ptr(dk),
        K._ptr(dv), B, Sq, H, Hkv, q.stride(1), q.stride(0), k.stride(1), k.stride(0), v.stride(1), v.stride(0),
        o.stride(1), o.stride(0), do.stride(1), do.stride(0), dq.stride(1), dq.stride(0), dk.stride(1), dk.stride(0),
        dv.stride(1), dv.stride(0), float(scale), 1 if causal else 0, K._stream())
    K._check(rc, "flash_attn_bwd")
    return dq, dk, dv


class _FlashAttn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, q, k, v, causal, scale):
        o, lse = flash_attention_forward(q, k, v, causal, scale)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.causal, ctx.scale = causal, scale
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        dq, dk, dv = flash_attention_backward(q, k, v, o, lse, do.contiguous(), ctx.causal, ctx.scale)
        return dq, dk, dv, None, None


class _FlashAttnLSE(torch.autograd.Function):
    """(out, lse) with gradients through both (partial-attention building block of chunked / blockwise schemes)."""

    @staticmethod
    def forward(ctx, q, k, v, causal, scale):
        o, lse = flash_attention_forward(q, k, v, causal, scale)
        ctx.save_for_backward(q, k, v, o, lse)
        ctx.causal, ctx.scale = causal, scale
        return o, lse

    @staticmethod
    def backward(ctx, do, dlse):
        q, k, v, o, lse = ctx.saved_tensors
        dq, dk, dv = flash_attention_backward(q, k, v, o, lse, do.contiguous(), ctx.causal, ctx.scale, dlse=dlse)
        return dq, dk, dv, None, None


def lse_attention_supported(q, k, v):
    """Can :func:`flash_attention_with_lse` serve these operands forward AND backward on the native kernels?"""
    return supported(q, k, v) and _native_backward_available() and q.shape[1] == k.shape[1] and q.shape[1] % 128 == 0


def flash_attention_with_lse(q, k, v, causal=True, scale=None):
    """-> (out [B,S,H,128] bf16, lse [B,H,S] fp32), differentiable through both outputs."""
    return _FlashAttnLSE.apply(q, k, v, causal, scale)


def flash_attention(q, k, v, causal=True, scale=None):
    """Differentiable attention on the native kernels. q [B,S,H,128], k/v [B,S,Hkv,128] -> [B,S,H,128]."""
    return _FlashAttn.apply(q, k, v, causal, scale)
