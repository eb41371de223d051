"""``DeepSpeedCPUAdam`` / ``DeepSpeedCPUAdagrad`` / ``DeepSpeedCPULion``: host optimizers for ZeRO-Offload.

Reference: ``ops/adam/cpu_adam.py`` (``DeepSpeedCPUAdam.step`` with optional fp16 param copy-back, ``rollback``),
``ops/adagrad/cpu_adagrad.py``, ``ops/lion/cpu_lion.py`` over ``csrc/adam/cpu_adam_impl.cpp``. The native side is
``csrc/cpu/cpu_optim.cpp`` (AVX-512 via ``-march=native`` auto-vectorisation + OpenMP tiles).
"""
import ctypes

import torch

from .cpu_lib import lib

_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


class _CPUOptimizer(torch.optim.Optimizer):
    supports_flat_step = True
    optimizer_id = 0

    def __init__(self, params, defaults):
        super().__init__(params, defaults)
        self.opt_id = _CPUOptimizer.optimizer_id
        _CPUOptimizer.optimizer_id += 1
        self._lib = lib()

    def _state_names(self):
        return ("exp_avg", "exp_avg_sq")

    def _init_state(self, p):
        st = self.state[p]
        if "step" not in st:
            st["step"] = 0
            for n in self._state_names():
                st[n] = torch.zeros_like(p, dtype=torch.float32, device="cpu")
        return st

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        loss = closure() if closure is not None else None
        grads, lp = {}, {}
        for gi, group in enumerate(self.param_groups):
            for pi, p in enumerate(group["params"]):
                if p.grad is not None:
                    grads[p] = p.grad
                    if fp16_param_groups is not None:
                        lp[p] = fp16_param_groups[gi][pi]
        self.step_flat(grads, lp, 1.0)
        return loss

    @torch.no_grad()
    def step_flat(self, grads, lp_params=None, grad_scale=1.0, advance_step=True):
        """``advance_step=False``: a further slice of the same optimizer step (sub-group stepping for NVMe swapping)."""
        lp_params = lp_params or {}
        for group in self.param_groups:
            used = [p for p in group["params"] if p in grads]
            if not used:
                continue
            if advance_step or "step" not in group:
                group["step"] = group.get("step", 0) + 1
            for p in used:
                assert p.device.type == "cpu" and p.dtype == torch.float32 and p.is_contiguous(), \
                    "CPU optimizers operate on contiguous fp32 host tensors"
                st = self._init_state(p)
                st["step"] = group["step"]
                g = grads[p]
                g = g if g.is_contiguous() else g.contiguous()
                lp = lp_params.get(p)
                host_lp = lp if (lp is not None and lp.device.type == "cpu") else None
                self._update(group, p, g, st, host_lp, grad_scale)
                if lp is not None and host_lp is None:
                    lp.copy_(p.to(lp.dtype), non_blocking=True)

    def _update(self, group, p, g, st, lp, grad_scale):
        raise NotImplementedError


class DeepSpeedCPUAdam(_CPUOptimizer):

    def __init__(self, model_params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0,
                 amsgrad=False, adamw_mode=True, fp32_optimizer_states=True):
        assert not amsgrad, "DeepSpeedCPUAdam does not support amsgrad"
        super().__init__(model_params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                            bias_correction=bias_correction))
        self.adam_w_mode = adamw_mode
        self.fp32_optimizer_states = fp32_optimizer_states

    def _update(self, group, p, g, st, lp, grad_scale):
        b1, b2 = group["betas"]
        self._lib.dsb200_cpu_adam(_p(p), _p(g), _DT[g.dtype], _p(st["exp_avg"]), _p(st["exp_avg_sq"]), _p(lp),
                                  _DT[lp.dtype] if lp is not None else 0, p.numel(), group["lr"], b1, b2, group["eps"],
                                  group["weight_decay"], st["step"], int(group["bias_correction"]),
                                  int(self.adam_w_mode), float(grad_scale))

    @torch.no_grad()
    def rollback(self, grads, grad_scale=1.0):
        """Undo the last step (speculative step rollback, reference ``ds_adam_rollback``)."""
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p not in grads:
                    continue
                st = self.state[p]
                g = grads[p].contiguous()
                self._lib.dsb200_cpu_adam_rollback(_p(p), _p(g), _DT[g.dtype], _p(st["exp_avg"]), _p(st["exp_avg_sq"]),
                                                   p.numel(), group["lr"], b1, b2, group["eps"], group["weight_decay"],
                                                   st["step"], int(group["bias_correction"]), int(self.adam_w_mode),
                                                   float(grad_scale))
                st["step"] -= 1
            group["step"] = group.get("step", 1) - 1


    # ---- sub-group form used by SuperOffload (reference ``cpu_adam.py:166-243``): the state lives under an integer
    # sub-group id instead of under the parameter, so the same flat fp32 buffer object can be stepped in slices ----------
    def _subgroup_state(self, subgroup_id, p):
        st = self.state[subgroup_id]
        if len(st) == 0:
            st["step"] = 0
            dt = torch.float if self.fp32_optimizer_states else p.dtype
            st["exp_avg"] = torch.zeros_like(p.data, dtype=dt, device="cpu")
            st["exp_avg_sq"] = torch.zeros_like(p.data, dtype=dt, device="cpu")
        return st

    @torch.no_grad()
    def step_subgroup(self, subgroup_id, closure=None):
        """One Adam step of the parameters that currently hold a gradient, with the state of ``subgroup_id``."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                assert p.device.type == "cpu", f"CPUAdam param is on {p.device} and must be 'cpu', make sure you " \
                    "enabled 'offload_optimizer': 'cpu' in your ZeRO config."
                st = self._subgroup_state(subgroup_id, p)
                st["step"] += 1
                g = p.grad.data.contiguous()
                self._update(group, p.data, g, st, None, 1.0)
        return loss

    @torch.no_grad()
    def rollback_subgroup(self, sub_group_id, closure=None):
        """Undo the last :meth:`step_subgroup` of ``sub_group_id`` (same gradients must still be attached)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if sub_group_id not in self.state or len(self.state[sub_group_id]) == 0:
            raise RuntimeError(f"Cannot rollback optimizer state for sub_group_id {sub_group_id} as it has not been "
                               "initialized.")
        st = self.state[sub_group_id]
        if st.get("step", 0) <= 0:
            raise RuntimeError(f"Cannot rollback sub_group_id {sub_group_id}: step count is {st.get('step', 0)}")
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                g = p.grad.data.contiguous()
                self._lib.dsb200_cpu_adam_rollback(_p(p.data), _p(g), _DT[g.dtype], _p(st["exp_avg"]),
                                                   _p(st["exp_avg_sq"]), p.numel(), group["lr"], b1, b2, group["eps"],
                                                   group["weight_decay"], st["step"], int(group["bias_correction"]),
                                                   int(self.adam_w_mode), 1.0)
                st["step"] -= 1
        return loss


class DeepSpeedCPUAdagrad(_CPUOptimizer):

    def __init__(self, model_params, lr=1e-2, eps=1e-10, weight_decay=0, amsgrad=False, fp32_optimizer_states=True):
        super().__init__(model_params, dict(lr=lr, eps=eps, weight_decay=weight_decay))

    def _state_names(self):
        return ("exp_avg_sq", )

    def _update(self, group, p, g, st, lp, grad_scale):
        self._lib.dsb200_cpu_adagrad(_p(p), _p(g), _DT[g.dtype], _p(st["exp_avg_sq"]), _p(lp),
                                     _DT[lp.dtype] if lp is not None else 0, p.numel(), group["lr"], group["eps"],
                                     group["weight_decay"], float(grad_scale))


class DeepSpeedCPULion(_CPUOptimizer):

    def __init__(self, model_params, lr=1e-3, betas=(0.9, 0.999), weight_decay=0, fp32_optimizer_states=True):
        super().__init__(model_params, dict(lr=lr, betas=betas, weight_decay=weight_decay))

    def _state_names(self):
        return ("exp_avg", )

    def _update(self, group, p, g, st, lp, grad_scale):
        b1, b2 = group["betas"]
        self._lib.dsb200_cpu_lion(_p(p), _p(g), _DT[g.dtype], _p(st["exp_avg"]), _p(lp),
                                  _DT[lp.dtype] if lp is not None else 0, p.numel(), group["lr"], b1, b2,
                                  group["weight_decay"], float(grad_scale))
