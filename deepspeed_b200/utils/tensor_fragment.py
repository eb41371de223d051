"""Debug/inspection API over sharded state: ``safe_get_full_fp32_param`` & friends.

Reference: ``utils/tensor_fragment.py:134-336`` (safe_get/set_full_{fp32_param,grad,optimizer_state} and the
``_local_`` variants). All ``full`` getters are collective (every rank of the shard group must call).
"""
from dataclasses import dataclass

import torch


@dataclass
class fragment_address:
    """(start, numel) of a parameter fragment inside a rank partition; stored in ``param_slice_mappings``."""
    numel: int
    start: int


fragment_address.__module__ = "deepspeed.utils.tensor_fragment"


def _zopt(param):
    u = getattr(param, "ds_unit", None)
    return getattr(u, "_zopt", None) if u is not None else None


def safe_get_full_fp32_param(param):
    z = _zopt(param)
    if z is None:
        return param.detach().float()
    return z.get_full_hp_param(param)


def safe_set_full_fp32_param(param, value):
    z = _zopt(param)
    if z is None:
        param.data.copy_(value.to(param.dtype))
    else:
        z.set_full_hp_param(value, param)


def safe_get_full_optimizer_state(param, optim_state_key):
    z = _zopt(param)
    return None if z is None else z.get_full_hp_param(param, optim_state_key=optim_state_key)


def safe_set_full_optimizer_state(param, value, optim_state_key):
    z = _zopt(param)
    if z is not None:
        z.set_full_hp_param(value, param, optim_state_key=optim_state_key)


def safe_get_full_grad(param):
    z = _zopt(param)
    if z is None:
        return None if param.grad is None else param.grad.detach().float()
    return z.get_full_grad(param)


def safe_set_full_grad(param, value):
    z = _zopt(param)
    if z is None:
        param.grad = value.to(param.dtype)
        return
    u, i = z._unit_of_param[id(param)]
    flat = value.reshape(-1)
    for (pi, s_in_p, s_in_shard, n) in u.param_fragments():
        if pi == i:
            u.grad_shard[s_in_shard:s_in_shard + n].copy_(flat[s_in_p:s_in_p + n].to(u.grad_shard.dtype))


def _local(param, tensor_of_unit):
    z = _zopt(param)
    if z is None:
        return None
    u, i = z._unit_of_param[id(param)]
    parts = [tensor_of_unit(u)[s:s + n] for (pi, _, s, n) in u.param_fragments() if pi == i]
    return torch.cat(parts) if parts else torch.empty(0)


def safe_get_local_fp32_param(param):
    return _local(param, lambda u: u.master_shard)


def safe_get_local_grad(param):
    return _local(param, lambda u: u.grad_shard)


def safe_get_local_optimizer_state(param, optim_state_key):
    z = _zopt(param)
    if z is None:
        return None
    return _local(param, lambda u: z._state_shard(u, optim_state_key))


def safe_set_local_fp32_param(param, value):
    z = _zopt(param)
    u, i = z._unit_of_param[id(param)]
    off = 0
    for (pi, _, s, n) in u.param_fragments():
        if pi == i:
            u.master_shard[s:s + n].copy_(value.reshape(-1)[off:off + n])
            off += n


def safe_set_local_grad(param, value):
    z = _zopt(param)
    u, i = z._unit_of_param[id(param)]
    off = 0
    for (pi, _, s, n) in u.param_fragments():
        if pi == i:
            u.grad_shard[s:s + n].copy_(value.reshape(-1)[off:off + n].to(u.grad_shard.dtype))
            off += n


def safe_update_full_grad_vectorized(params, update_fn):
    for p in params:
        g = safe_get_full_grad(p)
        if g is not None:
            safe_set_full_grad(p, update_fn(g, p))
