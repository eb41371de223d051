"""Fused optimizers: Adam/AdamW, Lion, LAMB, Adagrad, SGD front-ends over the sm_100a multi-tensor kernels.

Reference API studied: ``ops/adam/fused_adam.py`` (FusedAdam: ``adam_w_mode``, ``bias_correction``,
``set_grad_none``; kernel ``csrc/adam/multi_tensor_adam.cu:33-127``), ``ops/lion/fused_lion.py``,
``ops/lamb/fused_lamb.py`` (``csrc/lamb/fused_lamb_cuda_kernel.cu``: 3-phase trust ratio, ``max_coeff``/
``min_coeff``). B200 design: all tensors of a step are described by ONE device-resident chunk table (built once,
cached while pointers are stable) and updated by ONE launch; the ZeRO path additionally gets
``step_flat`` which consumes reduced grad shards directly (bf16 or fp32), applies unscale*clip, updates the fp32
master + moments and writes the low-precision shard in the same pass (reference needs 4 passes:
``stage_1_and_2.py:2350-2386``). On CPU (unit tests / plumbing config) the same math runs in PyTorch.
"""
import math

import torch

from . import kernels


def _on_cuda(t):
    return t.is_cuda and kernels.available()


class _FusedBase(torch.optim.Optimizer):
    supports_flat_step = True
    KIND = "adam"

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self._table_cache = {}

    def on_state_loaded(self):
        self._table_cache.clear()

    def zero_grad(self, set_to_none=True):
        for g in self.param_groups:
            for p in g["params"]:
                if set_to_none:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.zero_()

    # ---- state ---------------------------------------------------------------------------------
    def _state_names(self):
        return ("exp_avg", "exp_avg_sq")

    def _init_state(self, p, group):
        st = self.state[p]
        if "step" not in st:
            st["step"] = 0
            sdt = group.get("state_dtype", None) or p.dtype
            for n in self._state_names():
                st[n] = torch.zeros_like(p, dtype=sdt, memory_format=torch.preserve_format)
        return st

    # ---- generic step over p.grad ----------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None, grad_scaler=None):
        loss = closure() if closure is not None else None
        gmap, lp = {}, {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    gmap[p] = p.grad
        self.step_flat(gmap, lp, 1.0 if not scale else 1.0 / scale)
        return loss

    @torch.no_grad()
    def step_flat(self, grads, lp_params=None, grad_scale=1.0, advance_step=True):
        """``grads``: {param: grad tensor (any float dtype)}; ``lp_params``: {param: low-precision copy to write}.
        ``advance_step=False``: a further slice of the same optimizer step (the ZeRO wrapper steps unit by unit)."""
        lp_params = lp_params or {}
        items = []  # (group_idx, p, g, lp)
        for gi, group in enumerate(self.param_groups):
            any_p = False
            for p in group["params"]:
                g = grads.get(p)
                if g is None:
                    continue
                st = self._init_state(p, group)
                any_p = True
                items.append((gi, p, g, lp_params.get(p)))
            if any_p:
                if advance_step:
                    group["step"] = group.get("step", 0) + 1
                for p in group["params"]:
                    if p in grads:
                        self.state[p]["step"] = group.get("step", 1)
        if not items:
            return
        if all(_on_cuda(p) for _, p, _, _ in items):
            self._step_cuda(items, grad_scale)
        else:
            for gi, p, g, lp in items:
                self._step_torch(self.param_groups[gi], p, g.to(torch.float32) * grad_scale, lp)

    # ---- CUDA path -------------------------------------------------------------------------------
    def _hyper(self, group):
        raise NotImplementedError

    def _step_cuda(self, items, grad_scale):
        key = tuple((gi, p.data_ptr(), g.data_ptr(), (lp.data_ptr() if lp is not None else 0), p.numel())
                    for gi, p, g, lp in items)
        table = self._table_cache.get(key)
        if table is None:
            if len(self._table_cache) >= 512:  # one entry per (unit-sized) slice of the model; stale ones are dropped
                self._table_cache.clear()
            # kernel slots: [first moment (exp_avg / momentum_buffer) or None, second moment (exp_avg_sq) or None]
            states = [[next((self.state[p][n] for n in ("exp_avg", "momentum_buffer") if n in self.state[p]), None),
                       self.state[p].get("exp_avg_sq")] for _, p, _, _ in items]
            table = kernels.build_chunk_table([(gi, p, g, lp) for gi, p, g, lp in items], states)
            self._table_cache[key] = table
        hyper = [self._hyper(g) for g in self.param_groups]
        kernels.multi_tensor_optim(self.KIND, table, hyper, grad_scale)

    def _step_torch(self, group, p, g, lp):
        raise NotImplementedError


class FusedAdam(_FusedBase):
    """Adam / AdamW. ``adam_w_mode=True`` -> decoupled weight decay."""
    KIND = "adam"

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0., amsgrad=False, set_grad_none=True, state_dtype=None):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        state_dtype=state_dtype)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def _hyper(self, group):
        step = max(1, group.get("step", 1))
        b1, b2 = group["betas"]
        bc = group.get("bias_correction", True)
        bc1 = 1 - b1**step if bc else 1.0
        bc2 = 1 - b2**step if bc else 1.0
        return dict(lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"], wd=group["weight_decay"], bc1=bc1,
                    bc2=bc2, mode=self.adam_w_mode)

    def _step_torch(self, group, p, g, lp):
        st = self.state[p]
        h = self._hyper(group)
        m, v = st["exp_avg"], st["exp_avg_sq"]
        if h["mode"] == 0 and h["wd"] != 0:
            g = g.add(p.float(), alpha=h["wd"])
        m.mul_(h["beta1"]).add_(g.to(m.dtype), alpha=1 - h["beta1"])
        v.mul_(h["beta2"]).addcmul_(g.to(v.dtype), g.to(v.dtype), value=1 - h["beta2"])
        denom = (v.float() / h["bc2"]).sqrt_().add_(h["eps"])
        upd = (m.float() / h["bc1"]) / denom
        if h["mode"] == 1 and h["wd"] != 0:
            upd.add_(p.float(), alpha=h["wd"])
        p.add_(upd.to(p.dtype), alpha=-h["lr"])
        if lp is not None:
            lp.copy_(p)


class FusedLion(_FusedBase):
    KIND = "lion"

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0., set_grad_none=True, state_dtype=None):
        super().__init__(params, dict(lr=lr, betas=betas, weight_decay=weight_decay, state_dtype=state_dtype))
        self.set_grad_none = set_grad_none

    def _state_names(self):
        return ("exp_avg", )

    def _hyper(self, group):
        b1, b2 = group["betas"]
        return dict(lr=group["lr"], beta1=b1, beta2=b2, eps=0.0, wd=group["weight_decay"], bc1=1.0, bc2=1.0, mode=0)

    def _step_torch(self, group, p, g, lp):
        st = self.state[p]
        h = self._hyper(group)
        m = st["exp_avg"]
        if h["wd"] != 0:
            p.mul_(1 - h["lr"] * h["wd"])
        upd = (m.float() * h["beta1"] + g * (1 - h["beta1"])).sign_()
        p.add_(upd.to(p.dtype), alpha=-h["lr"])
        m.mul_(h["beta2"]).add_(g.to(m.dtype), alpha=1 - h["beta2"])
        if lp is not None:
            lp.copy_(p)


class FusedAdagrad(_FusedBase):
    KIND = "adagrad"

    def __init__(self, params, lr=1e-2, eps=1e-10, weight_decay=0., state_dtype=None):
        super().__init__(params, dict(lr=lr, eps=eps, weight_decay=weight_decay, state_dtype=state_dtype))

    def _state_names(self):
        return ("exp_avg_sq", )

    def _hyper(self, group):
        return dict(lr=group["lr"], beta1=0.0, beta2=0.0, eps=group["eps"], wd=group["weight_decay"], bc1=1.0,
                    bc2=1.0, mode=0)

    def _step_torch(self, group, p, g, lp):
        h = self._hyper(group)
        v = self.state[p]["exp_avg_sq"]
        if h["wd"] != 0:
            g = g.add(p.float(), alpha=h["wd"])
        v.addcmul_(g.to(v.dtype), g.to(v.dtype), value=1.0)
        p.addcdiv_(g.to(p.dtype), v.float().sqrt().add_(h["eps"]).to(p.dtype), value=-h["lr"])
        if lp is not None:
            lp.copy_(p)


class FusedSGD(_FusedBase):
    KIND = "sgd"

    def __init__(self, params, lr=1e-3, momentum=0.0, weight_decay=0., nesterov=False, state_dtype=None):
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay, nesterov=nesterov,
                                      state_dtype=state_dtype))

    def _state_names(self):
        return ("momentum_buffer", )

    def _hyper(self, group):
        return dict(lr=group["lr"], beta1=group["momentum"], beta2=1.0 if group["nesterov"] else 0.0, eps=0.0,
                    wd=group["weight_decay"], bc1=1.0, bc2=1.0, mode=1 if group.get("step", 1) <= 1 else 0)

    def _step_torch(self, group, p, g, lp):
        h = self._hyper(group)
        buf = self.state[p]["momentum_buffer"]
        if h["wd"] != 0:
            g = g.add(p.float(), alpha=h["wd"])
        if h["beta1"] != 0:
            if h["mode"] == 1:
                buf.copy_(g)
            else:
                buf.mul_(h["beta1"]).add_(g.to(buf.dtype))
            g = g.add(buf.float(), alpha=h["beta1"]) if h["beta2"] == 1.0 else buf.float()
        p.add_(g.to(p.dtype), alpha=-h["lr"])
        if lp is not None:
            lp.copy_(p)


class FusedLamb(_FusedBase):
    """LAMB with per-tensor trust ratio clamp [min_coeff, max_coeff] (reference ops/lamb/fused_lamb.py)."""
    KIND = "lamb"

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0., max_grad_norm=0., max_coeff=10.0, min_coeff=0.01, amsgrad=False, state_dtype=None):
        if amsgrad:
            raise RuntimeError("FusedLamb does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm, max_coeff=max_coeff,
                                      min_coeff=min_coeff, state_dtype=state_dtype))
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_coeffs = []

    def get_lamb_coeffs(self):
        return list(self.lamb_coeffs)

    def _hyper(self, group):
        step = max(1, group.get("step", 1))
        b1, b2 = group["betas"]
        bc = group.get("bias_correction", True)
        return dict(lr=group["lr"], beta1=b1, beta2=b2, eps=group["eps"], wd=group["weight_decay"],
                    bc1=(1 - b1**step) if bc else 1.0, bc2=(1 - b2**step) if bc else 1.0, mode=self.eps_mode,
                    max_coeff=group["max_coeff"], min_coeff=group["min_coeff"])

    def _lamb_update(self, group, p, g):
        st = self.state[p]
        h = self._hyper(group)
        m, v = st["exp_avg"], st["exp_avg_sq"]
        m.mul_(h["beta1"]).add_(g.to(m.dtype), alpha=1 - h["beta1"])
        v.mul_(h["beta2"]).addcmul_(g.to(v.dtype), g.to(v.dtype), value=1 - h["beta2"])
        mh, vh = m.float() / h["bc1"], v.float() / h["bc2"]
        denom = (vh + h["eps"]).sqrt() if h["mode"] == 0 else vh.sqrt() + h["eps"]
        upd = mh / denom + h["wd"] * p.float()
        return upd, h

    def _step_torch(self, group, p, g, lp):
        upd, h = self._lamb_update(group, p, g)
        wn, un = p.float().norm(), upd.norm()
        coeff = 1.0
        if wn > 0 and un > 0:
            coeff = float((wn / un).clamp(h["min_coeff"], h["max_coeff"]))
        self.lamb_coeffs.append(coeff)
        p.add_(upd.to(p.dtype), alpha=-h["lr"] * coeff)
        if lp is not None:
            lp.copy_(p)

    @torch.no_grad()
    def step_flat(self, grads, lp_params=None, grad_scale=1.0):
        self.lamb_coeffs = []
        # trust ratio is per *tensor*; under ZeRO a tensor is a shard fragment -> needs cross-rank norms.
        # Norm hooks can be installed by the ZeRO wrapper (``self.norm_all_reduce``); default is local.
        lp_params = lp_params or {}
        for gi, group in enumerate(self.param_groups):
            used = [p for p in group["params"] if p in grads]
            if not used:
                continue
            group["step"] = group.get("step", 0) + 1
            for p in used:
                self._init_state(p, group)
                self.state[p]["step"] = group["step"]
                g = grads[p].to(torch.float32) * grad_scale
                if _on_cuda(p):
                    kernels.lamb_step(p, g, self.state[p]["exp_avg"], self.state[p]["exp_avg_sq"], lp_params.get(p),
                                      self._hyper(group), self.lamb_coeffs)
                else:
                    self._step_torch(group, p, g, lp_params.get(p))


def multi_tensor_l2norm(tensors):
    """fp32 global L2 norm over a list of tensors (one launch on CUDA)."""
    if not tensors:
        return torch.zeros(())
    if all(_on_cuda(t) for t in tensors):
        return kernels.multi_tensor_sumsq(tensors).sqrt()
    return torch.sqrt(sum(t.float().pow(2).sum() for t in tensors))
