"""Communication-compressed optimizers: 1-bit Adam, 0/1 Adam, 1-bit LAMB.

Reference: ``runtime/fp16/onebit/adam.py:14`` (OnebitAdam), ``zoadam.py:14`` (ZeroOneAdam), ``lamb.py:15`` (OnebitLamb).
All three run a dense warm-up (ordinary gradient all-reduce by the engine), then switch the engine's gradient all-reduce
off and exchange a 1-bit, error-compensated quantity instead (:mod:`deepspeed_b200.runtime.comm.compressed`):

* 1-bit Adam: after ``freeze_step`` the variance is frozen and the *momentum* is what gets all-reduced.
* 0/1 Adam: the variance is refreshed at exponentially growing intervals (dense steps) and 1-bit gradients are used
  in between; after ``var_freeze_step`` workers take local steps and only synchronise the accumulated update every
  ``local_step_interval`` steps.
* 1-bit LAMB: as 1-bit Adam, with the layer-wise trust ratio frozen into per-tensor scaling coefficients learnt
  during warm-up and a bounded refresh factor afterwards.

One class hierarchy, state per parameter tensor; under the engine the parameters are the flat fp32 shards/replicas of
``ZeroShardedOptimizer`` (stage 0), so each "tensor" is large and the compressed collective is bandwidth efficient."""
import torch

from .. import comm as dist
from ..utils.logging import log_dist
from .comm.compressed import CompressedBackend


class _OnebitBase(torch.optim.Optimizer):
    compression_active = False  # read by ZeroShardedOptimizer: skip the dense gradient all-reduce when True

    def __init__(self, params, defaults, deepspeed=None, comm_backend_name="nccl", cuda_aware=False):
        super().__init__(params, defaults)
        self.deepspeed = deepspeed
        self.comm_backend_name = comm_backend_name
        self.using_pipeline = hasattr(deepspeed, "pipeline_enable_backward_allreduce")
        mpu = getattr(deepspeed, "mpu", None)
        self.comm_backend_handle = CompressedBackend(mpu) if dist.is_initialized() else None
        self.size = self.comm_backend_handle.size if self.comm_backend_handle is not None else 1

    def _set_dense_allreduce(self, on):
        """Dense gradient all-reduce by the engine on/off (reference: ``deepspeed.enable_backward_allreduce``)."""
        self.compression_active = not on
        if self.deepspeed is not None:
            if self.using_pipeline:
                self.deepspeed.pipeline_enable_backward_allreduce = on
            else:
                self.deepspeed.enable_backward_allreduce = on

    def _errors(self, state, p):
        if "worker_error" not in state:
            padded = self.comm_backend_handle.padded_size(p.numel()) if self.comm_backend_handle else p.numel()
            state["worker_error"] = torch.zeros(padded, device=p.device)
            state["server_error"] = torch.zeros(padded // self.size, device=p.device)
        return state["worker_error"], state["server_error"]

    def _compressed_mean(self, t, state, p):
        if self.size <= 1:
            return t
        we, se = self._errors(state, p)
        return self.comm_backend_handle.compressed_allreduce(t, we, se)

    def state_dict(self):
        sd = super().state_dict()
        for st in sd["state"].values():  # error-feedback buffers are not checkpointed (reference pops them on load)
            st.pop("worker_error", None)
            st.pop("server_error", None)
        return sd


class OnebitAdam(_OnebitBase):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm), deepspeed,
                         comm_backend_name, cuda_aware)
        self.freeze_step = freeze_step
        self.adam_freeze_key = False

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        last_step = 0
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                if grad.is_sparse:
                    raise RuntimeError("1-bit Adam does not support sparse gradients")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = 0
                    st["exp_avg"], st["exp_avg_sq"] = torch.zeros_like(p), torch.zeros_like(p)
                st["step"] += 1
                last_step = st["step"]
                m, v = st["exp_avg"], st["exp_avg_sq"]
                if not self.adam_freeze_key:
                    m.mul_(b1).add_(grad, alpha=1 - b1)
                    v.mul_(b2).addcmul_(grad, grad, value=1 - b2)
                elif group.get("non_freeze", False):
                    if self.size > 1:
                        dist.all_reduce(grad)
                        grad.mul_(1.0 / self.size)
                    m.mul_(b1).add_(grad, alpha=1 - b1)
                    v.mul_(b2).addcmul_(grad, grad, value=1 - b2)
                else:
                    m.mul_(b1).add_(grad, alpha=1 - b1)  # local gradient: the engine no longer all-reduces
                    m.copy_(self._compressed_mean(m, st, p))
                    if "exp_avg_mask" in group:
                        m.mul_(group["exp_avg_mask"].to(m.device))
                update = m / (v.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p
                p.add_(update, alpha=-group["lr"])
        if not self.adam_freeze_key and last_step >= self.freeze_step:
            log_dist("OnebitAdam - starting compressed communication", [0])
            self.adam_freeze_key = True
            self._set_dense_allreduce(False)
        return loss

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        steps = [st.get("step", 0) for st in self.state.values()]
        frozen = bool(steps) and max(int(s) for s in steps) >= self.freeze_step
        self.adam_freeze_key = frozen
        self._set_dense_allreduce(not frozen)
        for st in self.state.values():
            st.pop("worker_error", None)
            st.pop("server_error", None)


class ZeroOneAdam(_OnebitBase):

    def __init__(self, params, deepspeed=None, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., var_freeze_step=100000, var_update_scaler=16,
                 local_step_scaler=32678, local_step_clipper=16, amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("0/1 Adam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm), deepspeed,
                         comm_backend_name, cuda_aware)
        self.var_freeze_step = var_freeze_step
        self.var_update_scaler = var_update_scaler
        self.local_step_scaler = local_step_scaler
        self.local_step_clipper = local_step_clipper
        self.freeze_key = False

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        next_dense = None
        last_step = 0
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                st = self.state[p]
                if len(st) == 0:
                    st.update(step=0, exp_avg=torch.zeros_like(p), exp_avg_sq=torch.zeros_like(p), var_interval=1,
                              var_counter=0, local_step_interval=1, local_step_counter=0, lrs=0.0,
                              momentum_accumulator=torch.zeros_like(p))
                st["step"] += 1
                last_step = st["step"]
                m, v, acc = st["exp_avg"], st["exp_avg_sq"], st["momentum_accumulator"]
                denom_eps = group["eps"]
                if not self.freeze_key:
                    if st["step"] % st["var_interval"] == 0:  # dense step: the engine all-reduced this gradient
                        v.mul_(b2).addcmul_(grad, grad, value=1 - b2)
                        m.mul_(b1).add_(grad, alpha=1 - b1)
                    else:  # local gradient -> 1-bit mean
                        g1 = self._compressed_mean(grad.clone(), st, p)
                        if "exp_avg_mask" in group:
                            g1 = g1 * group["exp_avg_mask"].to(g1.device)
                        m.mul_(b1).add_(g1, alpha=1 - b1)
                else:
                    m.mul_(b1).add_(grad, alpha=1 - b1)
                    st["lrs"] += group["lr"]
                update = m / (v.sqrt() + denom_eps)
                if group["weight_decay"] > 0.0:
                    update = update + group["weight_decay"] * p
                p.add_(update, alpha=-group["lr"])
                if self.freeze_key:
                    acc.add_(update, alpha=-group["lr"])
                    if st["step"] % st["local_step_interval"] == 0:
                        # synchronise: undo the local displacement, average it in the preconditioned space
                        p.sub_(acc)
                        acc.mul_(v.sqrt() + denom_eps)
                        acc.copy_(self._compressed_mean(acc, st, p))
                        if "exp_avg_mask" in group:
                            acc.mul_(group["exp_avg_mask"].to(acc.device))
                        m.copy_(acc).div_(-st["lrs"])
                        p.add_(acc / (v.sqrt() + denom_eps))
                        acc.zero_()
                        st["lrs"] = 0.0
                    st["local_step_counter"] += 1
                    if st["local_step_counter"] == self.local_step_scaler:
                        st["local_step_counter"] = 0
                        st["local_step_interval"] = min(self.local_step_clipper, st["local_step_interval"] * 2)
                else:
                    if st["step"] % st["var_interval"] == 0:
                        st["var_counter"] += 1
                        if st["var_counter"] == self.var_update_scaler:
                            st["var_counter"] = 0
                            st["var_interval"] *= 2
                    next_dense = (st["step"] + 1) % st["var_interval"] == 0
        if not self.freeze_key:
            if last_step >= self.var_freeze_step:
                log_dist("ZeroOneAdam - variance frozen, starting local steps", [0])
                self.freeze_key = True
                self._set_dense_allreduce(False)
                # the error buffers held gradient-compression residuals; from now on they track the accumulated
                # update, a different quantity and scale
                for st in self.state.values():
                    if "worker_error" in st:
                        st["worker_error"].zero_()
                        st["server_error"].zero_()
            elif next_dense is not None:
                self._set_dense_allreduce(next_dense)
        return loss


class OnebitLamb(_OnebitBase):

    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., max_coeff=10.0, min_coeff=0.01,
                 amsgrad=False, cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0,
                 factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm, max_coeff=max_coeff,
                                      min_coeff=min_coeff), deepspeed, comm_backend_name, cuda_aware)
        self.freeze_step = freeze_step
        self.coeff_beta = coeff_beta
        self.factor_max, self.factor_min, self.factor_threshold = factor_max, factor_min, factor_threshold
        self.adam_freeze_key = False
        self.lamb_coeffs = []

    def get_lamb_coeffs(self):
        return self.lamb_coeffs

    @torch.no_grad()
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        self.lamb_coeffs = []
        last_step = 0
        for group in self.param_groups:
            b1, b2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                grad = p.grad
                st = self.state[p]
                if len(st) == 0:
                    st.update(step=0, exp_avg=torch.zeros_like(p), exp_avg_sq=torch.zeros_like(p),
                              lamb_coeff_freeze=0.0, last_factor=1.0)
                st["step"] += 1
                last_step = st["step"]
                m, v = st["exp_avg"], st["exp_avg_sq"]
                wd = group["weight_decay"]
                if not self.adam_freeze_key:  # warm-up: plain LAMB, remember a moving average of the trust ratio
                    m.mul_(b1).add_(grad, alpha=1 - b1)
                    v.mul_(b2).addcmul_(grad, grad, value=1 - b2)
                    if st["step"] == self.freeze_step:
                        st["exp_avg_sq_fresh"] = v.detach().clone()
                    update = m / (v.sqrt() + group["eps"])
                    if wd > 0.0:
                        update = update + wd * p
                    w_norm, u_norm = p.norm(), update.norm()
                    coeff = 1.0
                    if w_norm != 0 and u_norm != 0:
                        coeff = float((w_norm / u_norm).clamp(group["min_coeff"], group["max_coeff"]))
                    if coeff != 1.0:
                        st["lamb_coeff_freeze"] = self.coeff_beta * st["lamb_coeff_freeze"] + \
                            (1 - self.coeff_beta) * coeff
                    self.lamb_coeffs.append(coeff)
                    p.add_(update, alpha=-group["lr"] * coeff)
                else:  # compression stage: frozen variance, 1-bit momentum, frozen ratio x bounded refresh factor
                    if "scaling_coeff" not in st:
                        st["scaling_coeff"] = 1.0
                    m_last = st["exp_avg_last"] if "exp_avg_last" in st else m.detach().clone()  # same on all ranks
                    m.mul_(b1).add_(grad, alpha=1 - b1)
                    m.copy_(self._compressed_mean(m, st, p))
                    if "exp_avg_mask" in group:
                        m.mul_(group["exp_avg_mask"].to(m.device))
                    # estimate how the (uncommunicated) variance would have evolved from the reconstructed gradient
                    fresh = st.setdefault("exp_avg_sq_fresh", v.detach().clone())
                    grad_rec = (m - m_last * b1) / (1 - b1)  # gradient implied by the synchronised momentum
                    fresh.mul_(b2).addcmul_(grad_rec, grad_rec, value=1 - b2)
                    st["exp_avg_last"] = m.detach().clone()
                    denom = v.sqrt() + group["eps"]
                    update_prelim = m / denom
                    update = update_prelim + wd * p if wd > 0.0 else update_prelim
                    denom_real = fresh.sqrt() + group["eps"]
                    factor = float((denom / denom_real).max())
                    if wd > 0.0:
                        ratio = float(update_prelim.norm() / update.norm().clamp_min(1e-12))
                        factor = factor * ratio + (1.0 - ratio)
                    factor = min(max(factor, self.factor_min), self.factor_max)
                    lo, hi = st["last_factor"] * (1 - self.factor_threshold), st["last_factor"] * \
                        (1 + self.factor_threshold)
                    factor = min(max(factor, lo), hi)
                    st["last_factor"] = factor
                    coeff = st["lamb_coeff_freeze"] * factor
                    self.lamb_coeffs.append(coeff)
                    p.add_(update, alpha=-group["lr"] * coeff)
        if not self.adam_freeze_key and last_step >= self.freeze_step:
            log_dist("OnebitLamb - starting compressed communication", [0])
            self.adam_freeze_key = True
            self._set_dense_allreduce(False)
        return loss


def build_onebit_optimizer(name, model_parameters, params, engine):
    cls = {"onebitadam": OnebitAdam, "zerooneadam": ZeroOneAdam, "onebitlamb": OnebitLamb}[name.lower()]
    return cls(model_parameters, deepspeed=engine, **params)
