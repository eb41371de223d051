"""Helpers behind ``engine.offload_states`` (reference ``runtime/zero/offload_states.py:17-68``)."""
from typing import Set

import torch

from ...config.sections import OffloadStateTypeEnum  # noqa: F401


def _keys(optimizer):
    return [(p, k) for p, st in optimizer.state.items() for k, v in st.items() if torch.is_tensor(v) and v.numel() > 1]


def offload_adam_states(optimizer, device, pin_memory: bool = False, non_blocking: bool = False):
    """Move every tensor state (exp_avg, exp_avg_sq, ...) of ``optimizer`` to ``device``."""
    for p, k in _keys(optimizer):
        t = optimizer.state[p][k]
        if pin_memory and torch.device(device).type == "cpu" and torch.cuda.is_available():
            host = torch.empty_like(t, device="cpu").pin_memory()
            host.copy_(t, non_blocking=non_blocking)
            optimizer.state[p][k] = host
        else:
            optimizer.state[p][k] = t.to(device, non_blocking=non_blocking)


def reload_adam_states(optimizer, device, non_blocking: bool = False):
    for p, k in _keys(optimizer):
        optimizer.state[p][k] = optimizer.state[p][k].to(device, non_blocking=non_blocking)


offload_optimizer_states = offload_adam_states
reload_optimizer_states = reload_adam_states


def get_state_devices(model, state) -> Set[torch.device]:
    """Devices currently holding the given kind of state of an engine (used by the reference's tests)."""
    opt = model.optimizer
    name = getattr(state, "value", state)
    if name == "hp_params":
        return {st.master.device for st in opt.storages.values()}
    if name == "lp_grads":
        return {st.grad.device for st in opt.storages.values()}
    if name == "lp_params":
        return {(st.lp if st.lp is not None else st.master).device for st in opt.storages.values()}
    if name == "optim_states":
        return {v.device for st in opt.optimizer.state.values() for v in st.values()
                if torch.is_tensor(v) and v.numel() > 1}
    if name == "contiguous_grad_buffer":
        return {st.grad.device for st in opt.storages.values()}
    return set()
