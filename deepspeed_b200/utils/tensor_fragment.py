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


def safe_set_local_optimizer_state(param, value, optim_state_key):
    z = _zopt(param)
    u, i = z._unit_of_param[id(param)]
    dst = z._state_shard(u, optim_state_key)  # assembled copy of the unit's state shard
    if dst is None:
        return
    off = 0
    for (pi, _, s, n) in u.param_fragments():
        if pi == i:
            dst[s:s + n].copy_(value.reshape(-1)[off:off + n].to(dst.dtype))
            off += n
    z._write_state_shard(u, optim_state_key, dst)


# ---- method-style API: the reference binds these onto every low-precision parameter (``lp_param.get_full_hp_param()``,
# ``utils/mixed_precision_linkage.py:link_hp_params``); here they are free functions taking the parameter ------------
def get_full_hp_param(self, optim_state_key=None):
    if optim_state_key is None:
        return safe_get_full_fp32_param(self)
    return safe_get_full_optimizer_state(self, optim_state_key)


def set_full_hp_param(self, value, optim_state_key=None):
    if optim_state_key is None:
        safe_set_full_fp32_param(self, value)
    else:
        safe_set_full_optimizer_state(self, value, optim_state_key)


def get_full_hp_grad(self):
    return safe_get_full_grad(self)


def set_full_hp_grad(self, value):
    safe_set_full_grad(self, value)


class tensor_fragment:
    """View of the piece of one parameter that this rank's partition holds (reference ``tensor_fragment`` dataclass):
    built on demand from the unit bookkeeping instead of being stored per parameter."""

    def __init__(self, param):
        z = _zopt(param)
        assert z is not None, "parameter is not managed by a ZeRO optimizer"
        self.param = param
        self._z = z
        self._u, self._i = z._unit_of_param[id(param)]
        frags = [(s_in_p, s, n) for (pi, s_in_p, s, n) in self._u.param_fragments() if pi == self._i]
        self._frags = frags
        self.lp_fragment_address = fragment_address(numel=sum(n for _, _, n in frags),
                                                    start=frags[0][0] if frags else 0)
        self.hp_fragment_address = fragment_address(numel=sum(n for _, _, n in frags),
                                                    start=frags[0][1] if frags else 0)

    @property
    def hp_fragment(self):
        return safe_get_local_fp32_param(self.param)

    @property
    def lp_fragment(self):
        st = self._z.storages.get((self._u.group_key, self._u.dtype))
        lp = st.lp if st is not None and st.lp is not None else None
        if lp is None:
            return self.hp_fragment.to(self._u.dtype)
        return torch.cat([lp[self._u.shard_offset + s:self._u.shard_offset + s + n] for _, s, n in self._frags]) \
            if self._frags else torch.empty(0, dtype=self._u.dtype)

    def get_hp_fragment_address(self):
        return self.hp_fragment_address

    def get_optim_state_keys(self):
        keys = set()
        for st in self._z.optimizer.state.values():
            keys.update(k for k, v in st.items() if torch.is_tensor(v) and v.numel() > 1)
        return sorted(keys)

    def get_optim_state_fragment(self, key):
        out = safe_get_local_optimizer_state(self.param, key)
        if out is None:
            raise ValueError(f"{key} not found in optimizer state fragment")
        return out

    def get_hp_fragment(self, optim_state_key=None):
        return self.hp_fragment if optim_state_key is None else self.get_optim_state_fragment(optim_state_key)

    def update_hp(self):
        safe_set_local_fp32_param(self.param, self.lp_fragment.float())

    def update_lp(self):
        self._z.refresh_lp_from_master() if hasattr(self._z, "refresh_lp_from_master") else None


def get_hp_fragment_mapping(lp_param, *unused_args, **unused_kwargs):
    """The reference computes the overlap of a parameter with the rank's flat partition here; the unit bookkeeping
    already has it."""
    return tensor_fragment(lp_param)


def map_to_flat_opt_states(flat_hp_tensor, lp_tensors, optim_state, opt_keys):
    """Re-point the optimizer state of a flat master partition at one flat buffer per key, filled from the per-
    parameter fragments (reference ``tensor_fragment.py:72``; used after loading universal checkpoints)."""
    for key in opt_keys:
        hp = lp_tensors[0]._hp_mapping if lp_tensors and hasattr(lp_tensors[0], "_hp_mapping") else None
        dtype = hp.get_optim_state_fragment(key).dtype if hp is not None else flat_hp_tensor.dtype
        buffer = torch.zeros_like(flat_hp_tensor, dtype=dtype)
        for lp in lp_tensors:
            m = getattr(lp, "_hp_mapping", None)
            if m is None:
                continue
            frag = m.get_optim_state_fragment(key)
            addr = m.get_hp_fragment_address()
            buffer.narrow(0, addr.start, addr.numel).copy_(frag.reshape(-1))
        optim_state[flat_hp_tensor][key] = buffer
