"""Autograd front-ends of the fused sm_100a training ops: RMSNorm, LayerNorm, RoPE, SwiGLU, linear+cross-entropy.

On CUDA every function dispatches to ``libdsb200.so`` (raises if the library is missing); on CPU (plumbing tests)
the same math runs in plain PyTorch so models are device-agnostic.
"""
import math

import torch
import torch.nn.functional as F

from . import kernels as K


def _use_cuda(x):
    return x.is_cuda and K.available() and x.dtype in (torch.bfloat16, torch.float16)


# ---------------------------------------------------------------------------------------------------------
# RMSNorm (+ fused residual add)
# ---------------------------------------------------------------------------------------------------------
class _RMSNormFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, eps, residual):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        res2 = None
        if residual is not None:
            res2 = residual.reshape(-1, shape[-1])
            res2 = res2 if res2.is_contiguous() else res2.contiguous()
        y, res_out, rstd = K.rmsnorm_fwd(x2, res2, weight, eps)
        ctx.save_for_backward(res_out if residual is not None else x2, weight, rstd)
        ctx.has_res = residual is not None
        ctx.shape = shape
        if residual is not None:
            return y.view(shape), res_out.view(shape)
        return y.view(shape), None

    @staticmethod
    def backward(ctx, dy, dres):
        xin, weight, rstd = ctx.saved_tensors
        dy2 = dy.reshape(-1, ctx.shape[-1])
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        dres2 = None
        if ctx.has_res and dres is not None:
            dres2 = dres.reshape(-1, ctx.shape[-1])
            dres2 = dres2 if dres2.is_contiguous() else dres2.contiguous()
        dx, dw = K.rmsnorm_bwd(dy2, xin, weight, rstd, dres2)
        dx = dx.view(ctx.shape)
        return dx, dw, None, (dx if ctx.has_res else None)


def rms_norm(x, weight, eps=1e-6, residual=None):
    """``y = rmsnorm(x [+ residual]) * weight``. With ``residual`` returns ``(y, x + residual)``."""
    if _use_cuda(x):
        y, r = _RMSNormFn.apply(x, weight, eps, residual)
        return (y, r) if residual is not None else y
    if residual is not None:
        x = x + residual
    xf = x.float()
    y = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * weight
    return (y, x) if residual is not None else y


class _LayerNormFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        shape = x.shape
        x2 = x.reshape(-1, shape[-1])
        x2 = x2 if x2.is_contiguous() else x2.contiguous()
        y, stats = K.layernorm_fwd(x2, weight, bias, eps)
        ctx.save_for_backward(x2, weight, stats)
        ctx.shape, ctx.has_bias = shape, bias is not None
        return y.view(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, stats = ctx.saved_tensors
        dy2 = dy.reshape(-1, ctx.shape[-1])
        dy2 = dy2 if dy2.is_contiguous() else dy2.contiguous()
        dx, dw, db = K.layernorm_bwd(dy2, x2, weight, stats, ctx.has_bias)
        return dx.view(ctx.shape), dw, db, None


def layer_norm(x, weight, bias=None, eps=1e-5):
    if _use_cuda(x):
        return _LayerNormFn.apply(x, weight, bias, eps)
    return F.layer_norm(x, (x.shape[-1], ), weight, bias, eps)


# ---------------------------------------------------------------------------------------------------------
# RoPE
# ---------------------------------------------------------------------------------------------------------
def rope_tables(seq_len, dim, base=10000.0, device=None, scaling=None):
    """fp32 cos/sin tables [seq_len, dim/2]. ``scaling``: optional dict for Llama-3 style frequency scaling."""
    inv = 1.0 / (base**(torch.arange(0, dim, 2, dtype=torch.float32, device=device) / dim))
    if scaling and scaling.get("rope_type", scaling.get("type")) == "llama3":
        factor, lo, hi = scaling["factor"], scaling["low_freq_factor"], scaling["high_freq_factor"]
        old = scaling["original_max_position_embeddings"]
        wavelen = 2 * math.pi / inv
        smooth = ((old / wavelen) - lo) / (hi - lo)
        scaled = torch.where(wavelen > old / lo, inv / factor, inv)
        mid = (wavelen <= old / lo) & (wavelen >= old / hi)
        inv = torch.where(mid, (1 - smooth) * inv / factor + smooth * inv, scaled)
    t = torch.arange(seq_len, dtype=torch.float32, device=device)
    fr = torch.outer(t, inv)
    return fr.cos().contiguous(), fr.sin().contiguous()


class _RopeFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, cos, sin, pos, seq_len):
        ctx.save_for_backward(cos, sin, pos if pos is not None else torch.empty(0))
        ctx.has_pos, ctx.seq_len = pos is not None, seq_len
        out = x.clone() if not x.is_contiguous() else x.clone()
        K.rope_(out, cos, sin, pos, seq_len, backward=False)
        return out

    @staticmethod
    def backward(ctx, dout):
        cos, sin, pos = ctx.saved_tensors
        g = dout.contiguous().clone()
        K.rope_(g, cos, sin, pos if ctx.has_pos else None, ctx.seq_len, backward=True)
        return g, None, None, None, None


def _rope_torch(x, cos, sin, pos, seq_len):
    # x: [B, S, H, D]
    B, S, H, D = x.shape
    if pos is None:
        c, s = cos[:S], sin[:S]
        c, s = c.view(1, S, 1, D // 2), s.view(1, S, 1, D // 2)
    else:
        c, s = cos[pos.view(B, S)].view(B, S, 1, D // 2), sin[pos.view(B, S)].view(B, S, 1, D // 2)
    x1, x2 = x[..., :D // 2].float(), x[..., D // 2:].float()
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1).to(x.dtype)


def apply_rope(x, cos, sin, position_ids=None):
    """Half-rotation RoPE on ``x`` [B, S, H, D] (HF Llama convention). ``position_ids``: int [B, S] or None."""
    if _use_cuda(x):
        pos = position_ids.reshape(-1).to(torch.int32) if position_ids is not None else None
        return _RopeFn.apply(x, cos, sin, pos, x.shape[1])
    return _rope_torch(x, cos, sin, position_ids, x.shape[1])


class _RopeQKVFn(torch.autograd.Function):
    """In-place RoPE on the q and k heads of a fused QKV projection ``[B, S, (H + 2*Hkv) * D]`` (v untouched):
    one launch, no copies of q / k; backward rotates the incoming gradient in place the other way."""

    @staticmethod
    def forward(ctx, qkv, cos, sin, pos, n_rot_heads, n_heads_total, head_dim):
        B, S = qkv.shape[0], qkv.shape[1]
        ctx.save_for_backward(cos, sin, pos if pos is not None else torch.empty(0))
        ctx.has_pos, ctx.geom = pos is not None, (B, S, n_rot_heads, n_heads_total, head_dim)
        K.rope_strided_(qkv, cos, sin, pos, B * S, n_rot_heads, head_dim, n_heads_total * head_dim, S, backward=False)
        ctx.mark_dirty(qkv)
        return qkv

    @staticmethod
    def backward(ctx, dqkv):
        cos, sin, pos = ctx.saved_tensors
        B, S, n_rot, n_tot, hd = ctx.geom
        g = dqkv if dqkv.is_contiguous() else dqkv.contiguous()
        K.rope_strided_(g, cos, sin, pos if ctx.has_pos else None, B * S, n_rot, hd, n_tot * hd, S, backward=True)
        return g, None, None, None, None, None, None


def apply_rope_qkv(qkv, cos, sin, position_ids, n_q_heads, n_kv_heads, head_dim):
    """RoPE on the q and k parts of a fused projection output ``qkv`` [B, S, (Hq + 2*Hkv) * D] laid out
    ``[q heads | k heads | v heads]``. Returns ``(q [B,S,Hq,D], k [B,S,Hkv,D], v [B,S,Hkv,D])`` as views."""
    B, S = qkv.shape[0], qkv.shape[1]
    if _use_cuda(qkv) and qkv.is_contiguous():
        pos = position_ids.reshape(-1).to(torch.int32) if position_ids is not None else None
        qkv = _RopeQKVFn.apply(qkv, cos, sin, pos, n_q_heads + n_kv_heads, n_q_heads + 2 * n_kv_heads, head_dim)
        q, k, v = qkv.split([n_q_heads * head_dim, n_kv_heads * head_dim, n_kv_heads * head_dim], dim=-1)
        return (q.view(B, S, n_q_heads, head_dim), k.view(B, S, n_kv_heads, head_dim),
                v.view(B, S, n_kv_heads, head_dim))
    q, k, v = qkv.split([n_q_heads * head_dim, n_kv_heads * head_dim, n_kv_heads * head_dim], dim=-1)
    q = apply_rope(q.reshape(B, S, n_q_heads, head_dim), cos, sin, position_ids)
    k = apply_rope(k.reshape(B, S, n_kv_heads, head_dim), cos, sin, position_ids)
    return q, k, v.reshape(B, S, n_kv_heads, head_dim)


# ---------------------------------------------------------------------------------------------------------
# Linear whose weight gradient is written straight into the ZeRO flat gradient buffer
# ---------------------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    """``y = x W^T``. Backward computes ``dW = dy^T x`` with the GEMM output aimed at the parameter's region of the
    ZeRO unit's flat gradient buffer (``ZeroShardedOptimizer.grad_target``), so the gradient never exists as a
    separate tensor and the grad -> bucket copy of the hook path disappears (reference: the IPG-bucket copy
    ``stage3.py:1543-1566`` / ``stage_1_and_2.py:1193``). Falls back to returning ``dW`` to autograd."""

    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        # the parameter OBJECT (for the ZeRO bookkeeping): under non-reentrant activation checkpointing the saved tensors
        # come back as detached aliases without the parameter's attributes
        ctx.weight_param = weight
        return F.linear(x, weight)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.matmul(dy, weight)
        if ctx.needs_input_grad[1]:
            dy2, x2 = dy.reshape(-1, dy.shape[-1]), x.reshape(-1, x.shape[-1])
            weight = ctx.weight_param
            unit = getattr(weight, "ds_unit", None)
            zopt = getattr(unit, "_zopt", None) if unit is not None else None
            tgt = zopt.grad_target(weight, dy2.dtype) if zopt is not None and hasattr(zopt, "grad_target") else None
            if tgt is not None:
                view, acc = tgt
                view = view.view(dy2.shape[1], x2.shape[1])
                if acc:
                    view.addmm_(dy2.t(), x2)
                else:
                    torch.mm(dy2.t(), x2, out=view)
                zopt.grad_written(weight)
            else:
                dw = torch.mm(dy2.t(), x2)
        return dx, dw


def linear(x, weight, bias=None):
    if bias is None and weight.requires_grad and torch.is_grad_enabled() \
            and getattr(weight, "ds_unit", None) is not None:
        return _LinearFn.apply(x, weight)
    return F.linear(x, weight, bias)


class Linear(torch.nn.Linear):
    """``nn.Linear`` (same parameters / state dict) routed through :func:`linear`."""

    def forward(self, x):
        return linear(x, self.weight, self.bias)


# ---------------------------------------------------------------------------------------------------------
# SwiGLU
# ---------------------------------------------------------------------------------------------------------
class _SwiGLUFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, gate_up):
        ctx.save_for_backward(gate_up)
        return K.swiglu_fwd(gate_up)

    @staticmethod
    def backward(ctx, dy):
        (gate_up, ) = ctx.saved_tensors
        return K.swiglu_bwd(dy.contiguous(), gate_up)


def swiglu(gate_up):
    """``silu(gate) * up`` where ``gate_up = [..., 2*I]`` holds gate in the first half, up in the second."""
    if _use_cuda(gate_up) and gate_up.is_contiguous():
        return _SwiGLUFn.apply(gate_up)
    g, u = gate_up.chunk(2, dim=-1)
    return F.silu(g) * u


# ---------------------------------------------------------------------------------------------------------
# fused (chunked) linear + cross entropy: never materialises [tokens, vocab] in fp32
# ---------------------------------------------------------------------------------------------------------
class _LinearCrossEntropyFn(torch.autograd.Function):
    """loss = mean CE(h @ W^T, target). Gradients wrt h and W are produced during the forward pass, chunk by
    chunk (logits chunk -> in-place softmax-grad -> two GEMMs), so backward only rescales them."""

    @staticmethod
    def forward(ctx, h, weight, target, ignore_index, chunk_tokens, reduction_tokens):
        N, H = h.shape
        V = weight.shape[0]
        valid = (target != ignore_index)
        denom = reduction_tokens if reduction_tokens is not None else valid.sum().clamp(min=1)
        inv = 1.0 / denom if not torch.is_tensor(denom) else (1.0 / denom.float())
        need_h, need_w = h.requires_grad, weight.requires_grad
        dh = torch.empty_like(h) if need_h else None
        main_grad = getattr(weight, "_ds_grad_view", None)
        dw = None
        single = N <= chunk_tokens
        if need_w and main_grad is None and not single:
            dw = torch.zeros(V, H, dtype=torch.float32, device=h.device)
        loss_sum = torch.zeros((), dtype=torch.float32, device=h.device)
        inv_f = float(inv) if not torch.is_tensor(inv) else None
        inv_dev = inv.to(torch.float32).reshape(()).contiguous() if inv_f is None else None
        for s in range(0, N, chunk_tokens):
            e = min(N, s + chunk_tokens)
            logits = torch.matmul(h[s:e], weight.t())
            scale = inv_f if inv_f is not None else 1.0
            loss_rows = K.cross_entropy_inplace_grad(logits, target[s:e], ignore_index, scale,
                                                     None if inv_f is not None else inv_dev)
            loss_sum += loss_rows.sum()
            if need_h:
                torch.matmul(logits, weight, out=dh[s:e])
            if need_w:
                if single and main_grad is None:
                    dw = torch.matmul(logits.t(), h[s:e])
                elif dw is not None:
                    dw.add_(torch.matmul(logits.t(), h[s:e]))  # fp32 accumulation across chunks
                else:
                    main_grad.addmm_(logits.t(), h[s:e])
            del logits
        ctx.save_for_backward(dh if need_h else torch.empty(0), dw if dw is not None else torch.empty(0))
        ctx.need_h, ctx.need_w, ctx.w_dtype = need_h, need_w and dw is not None, weight.dtype
        return loss_sum * inv

    @staticmethod
    def backward(ctx, gloss):
        dh, dw = ctx.saved_tensors
        # the saved gradients are rescaled in place by a vectorised kernel that exits at once when the incoming gradient
        # is 1 (a broadcast multiply by a 0-dim device tensor takes the slow strided elementwise path: 2.1 ms for the
        # 128k x 4k lm_head gradient)
        if getattr(ctx, "_consumed", False):
            raise RuntimeError("linear_cross_entropy: backward called twice (gradients were rescaled in place)")
        ctx._consumed = True
        gh = gw = None
        g32 = gloss.to(torch.float32).reshape(()).contiguous()
        if ctx.need_h:
            gh = K.scale_by_device_scalar_(dh, g32) if dh.dtype != torch.float32 else dh.mul_(g32)
        if ctx.need_w:
            dw = K.scale_by_device_scalar_(dw, g32) if dw.dtype != torch.float32 else dw.mul_(g32)
            gw = dw if dw.dtype == ctx.w_dtype else dw.to(ctx.w_dtype)
        return gh, gw, None, None, None, None


def linear_cross_entropy(h, weight, target, ignore_index=-100, chunk_tokens=4096, reduction_tokens=None):
    """Mean token cross-entropy of ``h @ weight.T`` against ``target`` without keeping the logits."""
    h2 = h.reshape(-1, h.shape[-1])
    t2 = target.reshape(-1)
    if _use_cuda(h2):
        return _LinearCrossEntropyFn.apply(h2.contiguous(), weight, t2, ignore_index, chunk_tokens, reduction_tokens)
    logits = F.linear(h2, weight).float()
    if reduction_tokens is not None:
        return F.cross_entropy(logits, t2, ignore_index=ignore_index, reduction="sum") / reduction_tokens
    return F.cross_entropy(logits, t2, ignore_index=ignore_index)


class _CrossEntropyFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        loss_rows, lse = K.cross_entropy_fwd(logits, target, ignore_index)
        ctx.save_for_backward(logits, target, lse)
        ctx.ignore_index = ignore_index
        return loss_rows

    @staticmethod
    def backward(ctx, gl):
        logits, target, lse = ctx.saved_tensors
        p = torch.exp(logits.float() - lse.unsqueeze(1))
        valid = (target != ctx.ignore_index)
        p[torch.arange(p.shape[0], device=p.device)[valid], target[valid]] -= 1.0
        p = p * (gl * valid).unsqueeze(1)
        return p.to(logits.dtype), None, None


def cross_entropy_rows(logits, target, ignore_index=-100):
    """Per-row CE loss (fp32) for 16-bit logits [N, V]."""
    if _use_cuda(logits) and logits.is_contiguous():
        return _CrossEntropyFn.apply(logits, target, ignore_index)
    return F.cross_entropy(logits.float(), target, ignore_index=ignore_index, reduction="none")
