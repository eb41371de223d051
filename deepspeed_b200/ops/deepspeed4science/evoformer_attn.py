"""Evoformer attention for DS4Science / AlphaFold-style models (reference ``ops/deepspeed4science/evoformer_attn.py``
over the CUTLASS kernels in ``csrc/deepspeed4science/evoformer_attn``): attention over ``[*, L, H, D]`` tensors with
up to two additive biases - ``bias1 [B, N, 1, 1, L]`` (the MSA mask) and ``bias2 [B, 1, H, L, L]`` (the pair bias) -
and gradients for both.

Forward on a GPU: the tcgen05 flash-attention kernel with bias hooks (``csrc/attention_fwd.cu`` BIAS variant through
``ops/attention.py::flash_attention_bias_forward``; heads are zero-padded to 128, the MSA rows become the kernel's batch,
``bias2`` is shared by the N rows of a batch entry). The backward recomputes probabilities chunk by chunk with the
library GEMM from the saved log-sum-exp. CPU / unsupported dtypes: the chunked PyTorch forward below.

The score matrix is never materialised for the whole batch: rows of the MSA dimension are processed in chunks sized to
a fixed workspace, forward and backward (the backward recomputes each chunk's probabilities from the saved
log-sum-exp, like the fused kernel does)."""
import math

import torch

_WORKSPACE_ELEMS = 1 << 28  # fp32 score elements per chunk (1 GiB)


def _chunks(n_rows, per_row):
    step = max(1, _WORKSPACE_ELEMS // max(per_row, 1))
    return [(s, min(n_rows, s + step)) for s in range(0, n_rows, step)]


def _native_ok(q):
    import os
    if os.environ.get("DSB200_NATIVE_EVOFORMER", "1") != "1" or not q.is_cuda or q.shape[-1] > 128:
        return False
    if q.dtype not in (torch.bfloat16, torch.float16):
        return False  # the kernel computes in bf16: keep fp32 / fp64 inputs on the exact path
    from ..attention import bias_attention_supported
    return bias_attention_supported(q)


class EvoformerFusedAttention(torch.autograd.Function):

    @staticmethod
    def forward(ctx, q, k, v, bias1=None, bias2=None):
        # q, k, v: [B, N, L, H, D]
        B, N, L, H, D = q.shape
        scale = 1.0 / math.sqrt(D)
        out = torch.empty_like(q)
        acc = torch.float64 if q.dtype == torch.float64 else torch.float32
        lse = torch.empty(B, N, H, L, dtype=acc, device=q.device)
        native = _native_ok(q)
        if native:
            from ..attention import flash_attention_bias_forward
            o, l = flash_attention_bias_forward(
                q.reshape(B * N, L, H, D), k.reshape(B * N, L, H, D), v.reshape(B * N, L, H, D),
                bias1=None if bias1 is None else bias1.reshape(B * N, L),
                bias2=None if bias2 is None else bias2.reshape(B, H, L, L), bias2_div=N, scale=scale)
            out = o.reshape(B, N, L, H, D)
            lse = l.reshape(B, N, H, L).to(acc)
        for s, e in ([] if native else _chunks(N, B * H * L * L)):
            qc = q[:, s:e].permute(0, 1, 3, 2, 4)                    # [B, n, H, L, D]
            kc = k[:, s:e].permute(0, 1, 3, 2, 4)
            vc = v[:, s:e].permute(0, 1, 3, 2, 4)
            scores = torch.matmul(qc, kc.transpose(-1, -2)).to(acc) * scale
            if bias1 is not None:
                scores = scores + bias1[:, s:e].to(acc)              # [B, n, 1, 1, L]
            if bias2 is not None:
                scores = scores + bias2.to(acc)                      # [B, 1, H, L, L]
            l = torch.logsumexp(scores, dim=-1)
            lse[:, s:e] = l
            p = torch.exp(scores - l.unsqueeze(-1)).to(q.dtype)
            out[:, s:e] = torch.matmul(p, vc).permute(0, 1, 3, 2, 4)
        ctx.save_for_backward(q, k, v, out, lse, bias1 if bias1 is not None else torch.tensor([]),
                              bias2 if bias2 is not None else torch.tensor([]))
        ctx.has_b1, ctx.has_b2 = bias1 is not None, bias2 is not None
        ctx.native = native
        return out

    @staticmethod
    def backward(ctx, grad_output):
        q, k, v, out, lse, bias1, bias2 = ctx.saved_tensors
        B, N, L, H, D = q.shape
        scale = 1.0 / math.sqrt(D)
        acc = lse.dtype
        if ctx.native:
            return _backward_low_precision(ctx, grad_output, q, k, v, out, lse, bias1, bias2, scale)
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        need_b1 = ctx.has_b1 and ctx.needs_input_grad[3]
        need_b2 = ctx.has_b2 and ctx.needs_input_grad[4]
        db1 = torch.zeros(bias1.shape, dtype=acc, device=q.device) if need_b1 else None
        db2 = torch.zeros(bias2.shape, dtype=acc, device=q.device) if need_b2 else None
        delta = (grad_output.to(acc) * out.to(acc)).sum(-1).permute(0, 1, 3, 2)      # [B, N, H, L]
        for s, e in _chunks(N, B * H * L * L):
            qc = q[:, s:e].permute(0, 1, 3, 2, 4)
            kc = k[:, s:e].permute(0, 1, 3, 2, 4)
            vc = v[:, s:e].permute(0, 1, 3, 2, 4)
            doc = grad_output[:, s:e].permute(0, 1, 3, 2, 4)
            scores = torch.matmul(qc, kc.transpose(-1, -2)).to(acc) * scale
            if ctx.has_b1:
                scores = scores + bias1[:, s:e].to(acc)
            if ctx.has_b2:
                scores = scores + bias2.to(acc)
            p = torch.exp(scores - lse[:, s:e].unsqueeze(-1))                           # fp32 [B, n, H, L, L]
            dv[:, s:e] = torch.matmul(p.to(q.dtype).transpose(-1, -2), doc).permute(0, 1, 3, 2, 4)
            dp = torch.matmul(doc, vc.transpose(-1, -2)).to(acc)
            ds = p * (dp - delta[:, s:e].unsqueeze(-1))                                 # d(scores incl. biases)
            if need_b1:
                db1[:, s:e] += ds.sum(dim=(2, 3), keepdim=True)
            if need_b2:
                db2 += ds.sum(dim=1, keepdim=True)
            dsl = (ds * scale).to(q.dtype)
            dq[:, s:e] = torch.matmul(dsl, kc).permute(0, 1, 3, 2, 4)
            dk[:, s:e] = torch.matmul(dsl.transpose(-1, -2), qc).permute(0, 1, 3, 2, 4)
        return dq, dk, dv, (db1.to(grad_output.dtype) if need_b1 else None), \
            (db2.to(grad_output.dtype) if need_b2 else None)


def _backward_low_precision(ctx, grad_output, q, k, v, out, lse, bias1, bias2, scale):
    """Backward for the bf16 / fp16 GPU path: every GEMM in the input dtype, the probabilities recomputed from the saved
    log-sum-exp in fp32 one MSA chunk at a time, bias gradients accumulated in fp32. ``DSB200_EVO_BWD=sdpa`` instead
    differentiates the library's fused attention with the broadcast bias as mask."""
    import os
    B, N, L, H, D = q.shape
    need_b1 = ctx.has_b1 and ctx.needs_input_grad[3]
    need_b2 = ctx.has_b2 and ctx.needs_input_grad[4]
    if os.environ.get("DSB200_EVO_BWD", "gemm") == "sdpa":
        with torch.enable_grad():
            leaves = [t.detach().requires_grad_(True) for t in (q, k, v)]
            b1 = bias1.detach().requires_grad_(need_b1) if ctx.has_b1 else None
            b2 = bias2.detach().requires_grad_(need_b2) if ctx.has_b2 else None
            mask = None
            if b1 is not None or b2 is not None:
                mask = (b1 if b1 is not None else 0) + (b2 if b2 is not None else 0)
                mask = mask.expand(B, N, H, L, L)
            o = torch.nn.functional.scaled_dot_product_attention(
                leaves[0].transpose(2, 3), leaves[1].transpose(2, 3), leaves[2].transpose(2, 3), attn_mask=mask,
                scale=scale).transpose(2, 3)
            wanted = leaves + [t for t, need in ((b1, need_b1), (b2, need_b2)) if need]
            grads = list(torch.autograd.grad(o, wanted, grad_output))
        dq, dk, dv = grads[:3]
        rest = grads[3:]
        db1 = rest.pop(0) if need_b1 else None
        db2 = rest.pop(0) if need_b2 else None
        return dq, dk, dv, db1, db2
    # every pass over the [B, n, H, L, L] score tensor is in the input dtype and fused where the library has a fused
    # op (softmax, softmax backward); only the bias-gradient reductions accumulate in fp32
    dt = q.dtype
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    db1 = torch.zeros(B, N, L, dtype=torch.float32, device=q.device) if need_b1 else None
    db2 = torch.zeros(B, H, L, L, dtype=torch.float32, device=q.device) if need_b2 else None
    b2 = bias2.to(dt).reshape(B, 1, H, L, L) if ctx.has_b2 else None
    for s, e in _chunks(N, B * H * L * L):
        qc = (q[:, s:e] * scale).transpose(2, 3).contiguous()                                # [B, n, H, L, D]
        kc, vc, doc = (t[:, s:e].transpose(2, 3).contiguous() for t in (k, v, grad_output))
        x = torch.matmul(qc, kc.transpose(-1, -2))
        if b2 is not None:
            x.add_(b2)
        if ctx.has_b1:
            x.add_(bias1[:, s:e].to(dt))
        p = torch.softmax(x, dim=-1)
        del x
        dv[:, s:e] = torch.matmul(p.transpose(-1, -2), doc).transpose(2, 3)
        dp = torch.matmul(doc, vc.transpose(-1, -2))
        ds = torch._softmax_backward_data(dp, p, -1, dt)                                     # p * (dp - sum(dp * p))
        del dp, p
        if need_b1:
            db1[:, s:e] += ds.sum(dim=(2, 3), dtype=torch.float32)
        if need_b2:
            db2 += ds.sum(dim=1, dtype=torch.float32)
        dq[:, s:e] = torch.matmul(ds, kc).mul_(scale).transpose(2, 3)
        dk[:, s:e] = torch.matmul(ds.transpose(-1, -2), qc).transpose(2, 3)                   # qc carries the scale
    return dq, dk, dv, (db1.reshape(bias1.shape).to(bias1.dtype) if need_b1 else None), \
        (db2.reshape(bias2.shape).to(bias2.dtype) if need_b2 else None)


def DS4Sci_EvoformerAttention(Q, K, V, biases):
    """Q, K, V: [B, N, L, H, D]; ``biases``: up to two tensors, ``[B, N, 1, 1, L]`` and ``[B, 1, H, L, L]``."""
    assert len(biases) <= 2
    biases = list(biases) + [None] * (2 - len(biases))
    if biases[0] is not None:
        assert biases[0].shape == (Q.shape[0], Q.shape[1], 1, 1, Q.shape[2]), "bias1 shape is incorrect"
    if biases[1] is not None:
        assert biases[1].shape == (Q.shape[0], 1, Q.shape[3], Q.shape[2], Q.shape[2]), "bias2 shape is incorrect"
    return EvoformerFusedAttention.apply(Q, K, V, biases[0], biases[1])
