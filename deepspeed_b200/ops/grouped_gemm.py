"""Grouped GEMM for MoE experts: ``out[rows of expert e] = x[rows of e] @ w[e].T``.

Reference: Triton ``_group_gemm_kernel`` (``ops/triton_ops/group_gemm_triton.py:70-240``) on sm<90 and
``torch._grouped_mm`` on sm>=90 (``moe/ep_experts.py:92-135``). Here every expert's row block goes through the
hand-written tcgen05 GEMM (``csrc/gemm_tcgen05.cu``, K-major operands: exactly the ``x @ w[e].T`` form) in forward;
the backward uses the library GEMM for the two transposed products. CPU: plain matmul per expert.
"""
import torch

from . import kernels as K


def _use_native(x):
    import os
    return x.is_cuda and x.dtype == torch.bfloat16 and K.available() and os.environ.get("DSB200_NATIVE_GEMM", "1") == "1"


class _GroupedMM(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, w, counts):
        ctx.counts = counts
        ctx.w_param = w  # parameter object for the ZeRO bookkeeping (saved tensors may come back as detached aliases)
        ctx.save_for_backward(x, w)
        N = w.shape[1]
        out = x.new_empty(x.shape[0], N)
        off = 0
        native = _use_native(x) and x.shape[1] % 8 == 0 and N % 8 == 0
        for e, n in enumerate(counts):
            if n == 0:
                continue
            xe = x[off:off + n]
            if native:
                K.gemm_bf16_nt(xe, w[e], out[off:off + n])
            else:
                torch.matmul(xe, w[e].t(), out=out[off:off + n])
            off += n
        return out

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        g = g.contiguous()
        dx = torch.empty_like(x)
        # expert weight gradients go straight into the ZeRO flat gradient buffer when the optimizer offers a target
        # (ZeroShardedOptimizer.grad_target, same mechanism as ops.fused.Linear)
        wp = ctx.w_param
        unit = getattr(wp, "ds_unit", None)
        zopt = getattr(unit, "_zopt", None) if unit is not None else None
        tgt = zopt.grad_target(wp, g.dtype) if zopt is not None and hasattr(zopt, "grad_target") else None
        if tgt is not None:
            dw_view, acc = tgt[0].view(w.shape), tgt[1]
        else:
            dw_view, acc = torch.zeros_like(w), False
        off = 0
        for e, n in enumerate(ctx.counts):
            if n == 0:
                if tgt is not None and not acc:
                    dw_view[e].zero_()
                continue
            ge = g[off:off + n]
            torch.matmul(ge, w[e], out=dx[off:off + n])
            if acc:
                dw_view[e].addmm_(ge.t(), x[off:off + n])
            else:
                torch.matmul(ge.t(), x[off:off + n], out=dw_view[e])
            off += n
        if tgt is not None:
            zopt.grad_written(wp)
            return dx, None, None
        return dx, dw_view, None


def grouped_mm(x, w, counts):
    """x: [sum(counts), K] rows sorted by expert; w: [E, N, K]; counts: rows per expert (python ints)."""
    return _GroupedMM.apply(x, w, [int(c) for c in counts])
