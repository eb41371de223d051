"""``zero.Init`` (construct-time partitioning) and ``zero.GatheredParameters``.

Reference behaviour: ``runtime/zero/partition_parameters.py`` -- ``Init`` :940-1180 patches module
construction so every parameter is partitioned the moment its owning module finishes ``__init__``;
``_convert_to_deepspeed_param`` :1242 attaches ``ds_*`` attributes; ``GatheredParameters`` :2307 temporarily
re-assembles parameters (optionally broadcasting modifications made on ``modifier_rank``).

Design here: construction-time shards are *per parameter* (``p.ds_init_shard``); when the engine builds its
flat units (:mod:`.units`) each unit is assembled once from those shards (peak extra memory = one unit), after
which the unit layout is authoritative.
"""
import functools
import math
from typing import Iterable

import torch

from ... import comm as dist
from ...utils.logging import logger

_ACTIVE = []  # stack of active Init contexts
zero_init_context = 0


def _cdiv(a, b):
    return (a + b - 1) // b


def is_zero_param(p):
    return hasattr(p, "ds_id") or hasattr(p, "ds_init_shard") or hasattr(p, "ds_unit")


def gather_init_param(p):
    """Full tensor of a parameter created under ``zero.Init`` (collective over its partition group)."""
    shard = p.ds_init_shard
    group = getattr(p, "ds_process_group", None)
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world > 1:
        out = torch.empty(world * shard.numel(), dtype=shard.dtype, device=shard.device)
        src = shard.clone() if dist.get_backend(group) == "gloo" else shard
        dist.all_gather_into_tensor(out, src, group=group)
    else:
        out = shard
    return out[:p.ds_numel].view(p.ds_shape)


def _partition_param(p, group, rank, world, device, dtype):
    if hasattr(p, "ds_init_shard") or hasattr(p, "ds_unit"):
        return
    data = p.data
    if dtype is not None and data.is_floating_point():
        data = data.to(dtype)
    if device is not None and data.device != device:
        data = data.to(device)
    if world > 1:
        data = data.contiguous()
        dist.broadcast(data, dist.get_global_rank(group, 0), group=group)
    p.ds_shape = data.shape
    p.ds_numel = data.numel()
    p.ds_process_group = group
    per = _cdiv(max(p.ds_numel, 1), world)
    flat = data.reshape(-1)
    shard = torch.zeros(per, dtype=flat.dtype, device=flat.device)
    a, b = rank * per, min((rank + 1) * per, p.ds_numel)
    if b > a:
        shard[:b - a].copy_(flat[a:b])
    p.ds_init_shard = shard
    p.ds_tensor = shard
    p.ds_status = "NOT_AVAILABLE"
    p.data = torch.empty(0, dtype=flat.dtype, device=flat.device)
    Init.num_params += 1
    p.ds_id = Init.num_params


class Init:
    """Context manager: modules constructed inside have their parameters partitioned immediately."""
    num_params = 0

    def get_partition_dp_group(self, param):
        """The group ``param`` is partitioned over."""
        return getattr(param, "ds_process_group", self.group)

    def get_partition_rank(self):
        """Rank inside the partition group (subclasses may override, e.g. MiCS)."""
        return self.rank

    @property
    def num_partitions(self):
        return self.world

    def get_dp_process_group(self):
        """The communication group with all data-parallel ranks."""
        return self.group

    def __init__(self, module=None, data_parallel_group=None, mem_efficient_linear=True, remote_device=None,
                 pin_memory=False, config_dict_or_path=None, config=None, enabled=True, dtype=None, mpu=None,
                 zero_param_parallel_group=None, zero_quantized_weights=False,
                 zero_quantized_nontrainable_weights=False, sequence_data_parallel_group=None, param_swapper=None):
        self.enabled = enabled
        if config is not None and config_dict_or_path is None:
            config_dict_or_path = config
        self.dtype = dtype
        if config_dict_or_path is not None and dtype is None:
            from ...config import load_ds_config
            cd = load_ds_config(config_dict_or_path) if not hasattr(config_dict_or_path, "_param_dict") \
                else config_dict_or_path._param_dict
            if cd.get("bf16", {}).get("enabled", False):
                self.dtype = torch.bfloat16
            elif cd.get("fp16", {}).get("enabled", False):
                self.dtype = torch.float16
        if enabled and not dist.is_initialized():
            dist.init_distributed()
        self.group = sequence_data_parallel_group or data_parallel_group
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        self.device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        if remote_device == "cpu":
            self.device = torch.device("cpu")
        self._patched = []
        if module is not None and enabled:
            self._convert(module)

    # ---- patching ------------------------------------------------------------------------------
    def _wrap_init(self, cls):
        orig = cls.__dict__.get("__init__")
        if orig is None or getattr(orig, "_ds_wrapped", False):
            return
        ctx = self

        @functools.wraps(orig)
        def wrapper(module, *args, **kwargs):
            depth = getattr(module, "_ds_init_depth", 0)
            object.__setattr__(module, "_ds_init_depth", depth + 1)
            try:
                orig(module, *args, **kwargs)
            finally:
                object.__setattr__(module, "_ds_init_depth", depth)
            if depth == 0 and _ACTIVE and _ACTIVE[-1] is ctx:
                ctx._post_init(module)

        wrapper._ds_wrapped = True
        wrapper._ds_orig = orig
        cls.__init__ = wrapper
        self._patched.append(cls)

    def _all_subclasses(self, cls):
        seen, stack = set(), [cls]
        while stack:
            c = stack.pop()
            for s in c.__subclasses__():
                if s not in seen:
                    seen.add(s)
                    stack.append(s)
        return seen

    def __enter__(self):
        global zero_init_context
        if not self.enabled:
            return self
        _ACTIVE.append(self)
        zero_init_context += 1
        for cls in self._all_subclasses(torch.nn.Module):
            self._wrap_init(cls)
        self._orig_init_subclass = torch.nn.Module.__dict__.get("__init_subclass__")
        ctx = self

        def _init_subclass(cls, **kwargs):
            ctx._wrap_init(cls)

        torch.nn.Module.__init_subclass__ = classmethod(_init_subclass)
        return self

    def __exit__(self, exc_type, exc, tb):
        global zero_init_context
        if not self.enabled:
            return False
        for cls in self._patched:
            w = cls.__dict__.get("__init__")
            if w is not None and getattr(w, "_ds_wrapped", False):
                cls.__init__ = w._ds_orig
        self._patched = []
        if self._orig_init_subclass is None:
            try:
                del torch.nn.Module.__init_subclass__
            except AttributeError:
                pass
        else:
            torch.nn.Module.__init_subclass__ = self._orig_init_subclass
        _ACTIVE.pop()
        zero_init_context -= 1
        return False

    def _post_init(self, module):
        for p in module.parameters(recurse=True):
            _partition_param(p, self.group, self.rank, self.world, self.device, self.dtype)
        if self.device is not None:
            for name, buf in list(module.named_buffers(recurse=True)):
                if buf is not None and buf.device != self.device:
                    _set_buffer(module, name, buf.to(self.device))

    def _convert(self, module):
        for p in module.parameters(recurse=True):
            _partition_param(p, self.group, self.rank, self.world, self.device, self.dtype)


def _set_buffer(root, name, value):
    parts = name.split(".")
    m = root
    for k in parts[:-1]:
        m = getattr(m, k)
    m._buffers[parts[-1]] = value


def shutdown_init_context():
    """Temporarily disable active Init contexts (``deepspeed.initialize`` does this, reference __init__.py:163)."""
    for ctx in list(_ACTIVE):
        for cls in ctx._patched:
            w = cls.__dict__.get("__init__")
            if w is not None and getattr(w, "_ds_wrapped", False):
                cls.__init__ = w._ds_orig


def restore_init_context():
    for ctx in list(_ACTIVE):
        patched, ctx._patched = ctx._patched, []
        for cls in patched:
            ctx._wrap_init(cls)


class GatheredParameters:
    """Temporarily make the given (partitioned) parameters fully available.

    ``modifier_rank``: if not None, modifications made to the parameters by that rank inside the context are
    broadcast and written back to every shard on exit.
    """

    def __init__(self, params, modifier_rank=None, fwd_module=None, enabled=True):
        self.enabled = enabled
        if params is None:
            params = []
        elif isinstance(params, torch.nn.Parameter) or torch.is_tensor(params):
            params = [params]
        elif isinstance(params, torch.nn.Module):
            params = list(params.parameters())
        elif isinstance(params, Iterable):
            params = list(params)
        self.params = [p for p in params if is_zero_param(p)]
        self.modifier_rank = modifier_rank
        self.fwd_module = fwd_module
        self._units = []
        self._was_gathered = {}

    def __enter__(self):
        if not self.enabled:
            return
        seen = set()
        for p in self.params:
            if hasattr(p, "ds_unit"):
                u = p.ds_unit
                if u.uid in seen:
                    continue
                seen.add(u.uid)
                self._units.append(u)
                self._was_gathered[u.uid] = u.gathered
                z = _zopt_of(u)
                if z is not None and z.coordinator is not None and not u.persistent:
                    z.coordinator.gather(u, wait=True)
            elif hasattr(p, "ds_init_shard"):
                p.data = gather_init_param(p).clone()
                p.ds_status = "AVAILABLE"

    def __exit__(self, *exc):
        if not self.enabled:
            return False
        with torch.no_grad():
            written = []
            for u in self._units:
                z = _zopt_of(u)
                if self.modifier_rank is not None and u.full_param is not None:
                    if z is not None and z not in written:
                        # the shards are about to change under the peers' copy-engine gathers: wait for their reads,
                        # publish the new contents afterwards (ZeroShardedOptimizer._lp_shards_*)
                        if hasattr(z, "_join_step"):
                            z._join_step()
                        z._lp_shards_will_be_written()
                        written.append(z)
                    if u.world > 1:
                        src = dist.get_global_rank(u.dp_group, self.modifier_rank)
                        dist.broadcast(u.full_param, src, group=u.dp_group)
                    s, e = u.shard_range()
                    st = z.storages[(u.group_key, u.dtype)] if z is not None else None
                    if st is not None and st.lp is not None:
                        st.lp[u.shard_offset:u.shard_offset + u.shard_numel].copy_(u.full_param[s:e])
                    if u.param_shard is not None and u.param_shard.data_ptr() != u.full_param[s:e].data_ptr():
                        u.param_shard.copy_(u.full_param[s:e])
                    u.master_shard.copy_(u.full_param[s:e].to(u.master_shard.dtype))
                if z is not None and z.coordinator is not None and not u.persistent \
                        and not self._was_gathered.get(u.uid, False):
                    z.coordinator.release(u)
            for z in written:
                z._lp_shards_written()
            for p in self.params:
                if hasattr(p, "ds_init_shard") and not hasattr(p, "ds_unit"):
                    if self.modifier_rank is not None:
                        data = p.data.contiguous()
                        group = p.ds_process_group
                        world = dist.get_world_size(group) if dist.is_initialized() else 1
                        rank = dist.get_rank(group) if dist.is_initialized() else 0
                        if world > 1:
                            dist.broadcast(data, dist.get_global_rank(group, self.modifier_rank), group=group)
                        per = p.ds_init_shard.numel()
                        flat = data.reshape(-1)
                        a, b = rank * per, min((rank + 1) * per, p.ds_numel)
                        p.ds_init_shard.zero_()
                        if b > a:
                            p.ds_init_shard[:b - a].copy_(flat[a:b])
                    p.data = torch.empty(0, dtype=p.ds_init_shard.dtype, device=p.ds_init_shard.device)
                    p.ds_status = "NOT_AVAILABLE"
        return False


def _zopt_of(u):
    return getattr(u, "_zopt", None)


def register_external_parameter(module, parameter):
    """Declare that ``module.forward`` uses ``parameter`` owned by another module."""
    if hasattr(parameter, "ds_unit"):
        z = _zopt_of(parameter.ds_unit)
        if z is not None and z.coordinator is not None:
            z.coordinator.register_external_parameter(module, parameter)
            return
    module.__dict__.setdefault("_ds_external_parameters", []).append(parameter)


def unregister_external_parameter(module, parameter):
    lst = module.__dict__.get("_ds_external_parameters", [])
    if parameter in lst:
        lst.remove(parameter)


# names user code imports from the reference's ``runtime.zero.partition_parameters`` module
from .param_status import DeepSpeedTensorOverride, ZeroParamStatus, ZeroParamType  # noqa: E402,F401
