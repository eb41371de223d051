"""Runtime helpers (reference ``runtime/utils.py``): grad-norm / clipping over DP(+MP) groups, overflow check,
balanced partitioning, DummyOptim, see_memory_usage re-export."""
import math
from bisect import bisect_left

import torch

from .. import comm as dist
from ..utils.memory import see_memory_usage  # noqa: F401


class DummyOptim:
    """Placeholder optimizer when the client passes none (reference ``runtime/utils.py:39``)."""

    def __init__(self, params):
        self.param_groups = [{"params": params}]


def noop_decorator(func):
    return func


def is_model_parallel_parameter(p):
    return getattr(p, "model_parallel", False) or getattr(p, "tensor_model_parallel", False)


def get_global_norm(norm_list):
    return math.sqrt(sum(n**2 for n in norm_list))


def get_global_norm_of_tensors(input_tensors, norm_type=2, mpu=None, use_graph=False, moe_ep_group=None):
    """Norm over a list of tensors, reduced over the model-parallel (and expert-parallel) groups."""
    tensors = [t for t in input_tensors if t is not None]
    if not tensors:
        return 0.0
    dev = tensors[0].device
    if norm_type == math.inf:
        total = torch.stack([t.detach().abs().max().float() for t in tensors]).max()
        for g in (mpu.get_model_parallel_group() if mpu is not None else None, moe_ep_group):
            if g is not None:
                dist.all_reduce(total, op=dist.ReduceOp.MAX, group=g)
        return float(total)
    total = torch.zeros((), dtype=torch.float32, device=dev)
    for t in tensors:
        total += t.detach().float().norm(norm_type)**norm_type
    for g in (mpu.get_model_parallel_group() if mpu is not None else None, moe_ep_group):
        if g is not None:
            dist.all_reduce(total, group=g)
    out = float(total)**(1.0 / norm_type)
    return -1.0 if (math.isinf(out) or math.isnan(out)) else out


def clip_grad_norm_(parameters, max_norm, norm_type=2, mpu=None):
    """Megatron-aware ``clip_grad_norm_``: tensor-parallel duplicates are counted once."""
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = [p for p in parameters if p.grad is not None]
    mp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    grads = [p.grad for p in parameters if mp_rank == 0 or is_model_parallel_parameter(p)]
    total_norm = get_global_norm_of_tensors(grads, norm_type, mpu)
    clip_coef = max_norm / (total_norm + 1e-6)
    if clip_coef < 1:
        for p in parameters:
            p.grad.detach().mul_(clip_coef)
    return total_norm


def clip_tensors_by_global_norm(input_tensors, max_norm=1.0, global_norm=None, mpu=None, eps=1e-6, use_graph=False):
    if global_norm is None:
        global_norm = get_global_norm_of_tensors(input_tensors, mpu=mpu)
    clip_coef = max_norm / (global_norm + eps)
    if clip_coef < 1:
        for t in input_tensors:
            t.detach().mul_(clip_coef)
    return global_norm


def mask_nan_or_inf_with_val_inplace(input, device=None, val=-1.0):
    input.masked_fill_(~torch.isfinite(input), val)


class CheckOverflow:
    """inf/nan detection over parameter gradients with a MAX all-reduce (reference ``runtime/utils.py`` CheckOverflow)."""

    def __init__(self, param_groups=None, mpu=None, zero_reduce_scatter=False, deepspeed=None):
        self.mpu, self.params = mpu, []
        self.zero_reduce_scatter, self.deepspeed = zero_reduce_scatter, deepspeed
        for group in (param_groups or []):
            self.params.extend(group)

    @staticmethod
    def _has_inf_or_nan(x, i=None):
        s = float(x.float().sum())
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    def has_overflow(self, params, has_moe_params=None):
        overflow = self.has_overflow_serial(params)
        t = torch.tensor([float(overflow)], device=params[0].device if params else "cpu")
        if dist.is_initialized():
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if self.mpu is not None:
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.mpu.get_model_parallel_group())
        return bool(t.item())

    def check(self, param_groups=None):
        params = [p for g in (param_groups or [self.params]) for p in g]
        return self.has_overflow(params)


def partition_uniform(num_items, num_parts):
    from .pipe.module import partition_uniform as _pu
    return _pu(num_items, num_parts)


def partition_balanced(weights, num_parts):
    from .pipe.module import partition_balanced as _pb
    return _pb(weights, num_parts)


def prefix_sum_inc(weights):
    out, s = [], 0
    for w in weights:
        s += w
        out.append(s)
    return out


def all_gather_dp_groups(groups_flat, partitioned_param_groups, dp_process_group, start_alignment_factor=1,
                         allgather_bucket_size=None):
    """Reference ``runtime/utils.py:1016``: in-place all-gather of every flat param group from its partitions."""
    for flat, parts, pg in zip(groups_flat, partitioned_param_groups, dp_process_group):
        rank = dist.get_rank(pg)
        dist.all_gather_into_tensor(flat, parts[rank].contiguous(), group=pg)


def get_only_unique_item(items):
    s = set(items)
    if len(s) != 1:
        raise RuntimeError(f"expected there to be only one unique element in {items}")
    return next(iter(s))


def align_dense_tensors(tensor_list, alignment):
    num = sum(t.numel() for t in tensor_list)
    rem = num % alignment
    if rem:
        pad = torch.zeros(alignment - rem, device=tensor_list[0].device, dtype=tensor_list[0].dtype)
        return tensor_list + [pad]
    return tensor_list


def required_torch_version(min_version=None, max_version=None):
    from packaging import version
    v = version.parse(torch.__version__.split("+")[0])
    if min_version is not None and v < version.parse(str(min_version)):
        return False
    if max_version is not None and v > version.parse(str(max_version)):
        return False
    return True


def call_to_str(base, *args, **kwargs):
    name = f"{base}("
    if args:
        name += ", ".join(repr(a) for a in args)
        if kwargs:
            name += ", "
    if kwargs:
        name += ", ".join(f"{k}={v!r}" for k, v in kwargs.items())
    return name + ")"


# =====================================================================================================
# The rest of the reference module's helper surface (``runtime/utils.py``): seeding, device moves, norms of
# weights / flattened gradients, MoE-aware norms, partitioned tensors for pipeline activation partitioning, output
# backward hooks, memory reports, CUDA-graph helper, Adam state offload.
# =====================================================================================================
import contextlib  # noqa: E402
import gc  # noqa: E402
import os  # noqa: E402
from math import inf  # noqa: E402

PIPE_REPLICATED = "ds_pipe_replicated"


def set_random_seed(seed):
    """Seed python, numpy and torch (numpy takes 32-bit seeds, torch 64-bit)."""
    import random

    import numpy
    random.seed(seed)
    numpy.random.seed(seed % (2**32))
    torch.manual_seed(seed % (2**63))


class noop_context:

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        return False


def ensure_directory_exists(filename):
    os.makedirs(os.path.dirname(filename), exist_ok=True)


def _map_structure(item, fn):
    if torch.is_tensor(item):
        return fn(item)
    if isinstance(item, list):
        return [_map_structure(v, fn) for v in item]
    if isinstance(item, tuple):
        return tuple(_map_structure(v, fn) for v in item)
    if isinstance(item, dict):
        return {k: _map_structure(v, fn) for k, v in item.items()}
    return item


def copy_to_device(item, device, criterion_func):
    """Copy (clone) the tensors of a nested structure for which ``criterion_func(tensor)`` holds to ``device``."""
    return _map_structure(item, lambda t: t.to(device).clone() if criterion_func(t) else t)


def move_to_device(item, device, criterion_func=None):
    """Move (in place, through ``.data``) the selected tensors of a nested structure to ``device``."""

    def mv(t):
        if criterion_func is None or criterion_func(t):
            t.data = t.data.to(device)
        return t

    return _map_structure(item, mv)


def get_tensor_bytes(item):
    total = [0]
    _map_structure(item, lambda t: total.__setitem__(0, total[0] + t.numel() * t.element_size()) or t)
    return total[0]


def filter_empty_parameters(params):
    """Drop parameters with ``numel() == 0`` (optimizers that divide by numel would produce NaNs); accepts a list of
    parameters or of param-group dicts and returns the same form."""
    if not isinstance(params, (list, tuple)) or len(params) == 0:
        return params
    if isinstance(params[0], dict):
        out = []
        for group in params:
            g = dict(group)
            g["params"] = [p for p in group["params"] if p.numel() > 0]
            if g["params"]:
                out.append(g)
        return out
    return [p for p in params if p.numel() > 0]


def _mp_group(mpu):
    return mpu.get_model_parallel_group() if mpu is not None else None


def _reduced_norm(local, norm_type, mpu):
    """local: sum of |x|^p (or the max for inf) of this rank -> global norm over the model-parallel group."""
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    t = torch.tensor([float(local)], dtype=torch.float32, device=dev)
    if mpu is not None and dist.is_initialized():
        dist.all_reduce(t, op=dist.ReduceOp.MAX if norm_type == inf else dist.ReduceOp.SUM, group=_mp_group(mpu))
    total = float(t[0])
    if norm_type != inf:
        total = total**(1.0 / norm_type)
    if total in (float("inf"), -float("inf")) or total != total:
        return -1
    return total


def _counts_on_this_rank(p, mpu):
    """Tensor-parallel replicas of a non-sharded parameter are counted on tp rank 0 only; pipeline replicas once."""
    from ..utils.bwc import bwc_tensor_model_parallel_rank
    if getattr(p, PIPE_REPLICATED, False):
        return False
    return bwc_tensor_model_parallel_rank(mpu) == 0 or is_model_parallel_parameter(p)


def get_weight_norm(parameters, norm_type=2, mpu=None):
    """Norm of the parameters viewed as one vector, over the model-parallel group; -1 for inf / nan."""
    if torch.is_tensor(parameters):
        parameters = [parameters]
    norm_type = float(norm_type)
    if norm_type == inf:
        return _reduced_norm(max(float(p.data.abs().max()) for p in parameters), inf, mpu)
    local = sum(float(p.data.float().norm(norm_type))**norm_type for p in parameters if _counts_on_this_rank(p, mpu))
    return _reduced_norm(local, norm_type, mpu)


def get_flattened_grad_norm(parameters, norm_type=2, mpu=None, grad_norm_mask=None):
    """Gradient norm of flattened parameter groups. ``grad_norm_mask[i]``: ``[n, 2]`` tensor of ``[start, end)`` ranges
    of group ``i`` to leave out (regions owned by another tensor-parallel rank)."""
    if torch.is_tensor(parameters):
        parameters = [parameters]
    items = [(i, p) for i, p in enumerate(parameters) if p.grad is not None]
    norm_type = float(norm_type)
    if norm_type == inf:
        return _reduced_norm(max(float(p.grad.data.abs().max()) for _, p in items), inf, mpu)
    local = 0.0
    for i, p in items:
        g = p.grad.data.float()
        if grad_norm_mask is not None and len(grad_norm_mask[i]) > 0:
            keep = torch.ones_like(g, dtype=torch.bool)
            for a, b in grad_norm_mask[i].tolist():
                keep.view(-1)[int(a):int(b)] = False
            g = g * keep
        local += float(g.norm(norm_type))**norm_type
    return _reduced_norm(local, norm_type, mpu)


def get_grad_zeros(parameters, mpu=None):
    """Number of exactly-zero gradient elements over the model-parallel group."""
    if torch.is_tensor(parameters):
        parameters = [parameters]
    local = sum(float(p.grad.numel() - torch.count_nonzero(p.grad)) for p in parameters
                if p.grad is not None and _counts_on_this_rank(p, mpu))
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    t = torch.tensor([local], dtype=torch.float32, device=dev)
    if mpu is not None and dist.is_initialized():
        dist.all_reduce(t, group=_mp_group(mpu))
    total = float(t[0])
    return -1 if total in (float("inf"), -float("inf")) or total != total else total


def get_inactive_params(param_list):
    return [p for p in param_list if hasattr(p, "ds_id") and getattr(p, "ds_status", None) == "NOT_AVAILABLE"]


def get_norm_with_moe_layers_fast(all_groups_norm, group):
    """Average the norm over ``group``: expert gradients are not synchronised by the dense all-reduce, so ranks would
    otherwise clip with different norms."""
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    t = torch.tensor(all_groups_norm / float(dist.get_world_size(group=group)), dtype=torch.float, device=dev)
    dist.all_reduce(t, group=group)
    return t.item()


def get_norm_with_moe_layers(non_expert_norm, mpu, expert_tensors, norm_type=2):
    """Global norm from the dense norm plus, per expert group name, the norm of that group's gradient tensors reduced
    over its expert-parallel group. -1 when not finite."""
    from ..utils import groups
    norms = [float(non_expert_norm)]
    for name, tensors in expert_tensors.items():
        norms.append(float(get_global_norm_of_tensors(tensors, norm_type=norm_type, mpu=mpu, use_graph=False,
                                                      moe_ep_group=groups._get_expert_parallel_group(name))))
    if norm_type == inf:
        total = max(norms)
    else:
        total = sum(n**norm_type for n in norms)**(1.0 / norm_type)
    return -1 if total in (float("inf"), -float("inf")) or total != total else total


def all_gather_into_tensor_dp_groups(groups_flat, partitioned_param_groups, dp_process_group):
    for gid, (flat, parts) in enumerate(zip(groups_flat, partitioned_param_groups)):
        g = dp_process_group[gid]
        if dist.get_world_size(group=g) == 1:
            continue
        dist.all_gather_into_tensor(flat, parts[dist.get_rank(group=g)], group=g)


class PartitionedTensor:
    """A tensor split evenly (by elements) over a group, each rank holding one slice: used to partition pipeline
    activations over the tensor-parallel ranks. ``to_meta()`` + ``data()`` serialise it for p2p transfer,
    ``from_meta`` rebuilds it on the receiver, ``full()`` all-gathers the original."""

    def __init__(self, tensor, group, partition_meta=None):
        self.group = group
        self.num_parts = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        self.orig_size = list(tensor.size())
        self.orig_device = tensor.device
        self.partition = partition_uniform(num_items=tensor.numel(), num_parts=self.num_parts)
        a, b = self.partition[self.rank], self.partition[self.rank + 1]
        self.local_data = tensor.detach().contiguous().view(-1)[a:b].clone()
        self.even_split = tensor.numel() % self.num_parts == 0

    @classmethod
    def from_meta(cls, meta, local_part, group, device=None):
        assert meta.dtype == torch.long
        obj = cls(tensor=torch.ones(dist.get_world_size(group=group)), group=group)
        m = meta.tolist()
        ndim = m[0]
        obj.orig_size = m[1:1 + ndim]
        rest = m[1 + ndim:]
        assert obj.num_parts == rest[0] and obj.rank == rest[1]
        obj.partition = rest[2:]
        obj.orig_device = device if device is not None else local_part.device
        obj.local_data = local_part.detach()
        numel = 1
        for s in obj.orig_size:
            numel *= s
        obj.even_split = numel % obj.num_parts == 0
        return obj

    def to_meta(self):
        meta = [len(self.orig_size)] + list(self.orig_size) + [self.num_parts, self.rank] + list(self.partition)
        return torch.tensor(meta, dtype=torch.long, device=self.orig_device)

    def data(self):
        return self.local_data

    def local_size(self):
        return self.local_data.size()

    def full_size(self):
        return self.orig_size

    def full(self, device=None):
        device = device if device is not None else self.orig_device
        numel = 1
        for s in self.orig_size:
            numel *= s
        flat = torch.zeros(numel, dtype=self.local_data.dtype, device=device)
        if self.even_split:
            dist.all_gather_into_tensor(flat, self.local_data.to(device), group=self.group)
        else:
            for part in range(self.num_parts):
                buf = flat[self.partition[part]:self.partition[part + 1]]
                if part == self.rank:
                    buf.copy_(self.local_data)
                dist.broadcast(buf, dist.get_global_rank(self.group, part), group=self.group)
        return flat.view(self.orig_size).clone().detach()


_mem = {"alloc": 0, "cached": 0}


def memory_status(msg, print_rank=-1, reset_max=False):
    """Print allocated / reserved device memory and the change since the previous call."""
    rank = dist.get_rank() if dist.is_initialized() else 0
    if print_rank != -1 and rank != print_rank:
        return
    if not torch.cuda.is_available():
        print(f"RANK={rank} MEMSTATS {msg} (no accelerator)")
        return
    torch.cuda.synchronize()
    if reset_max:
        torch.cuda.reset_peak_memory_stats()
    alloc, cached = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    d_alloc, d_cached = alloc - _mem["alloc"], cached - _mem["cached"]
    _mem["alloc"], _mem["cached"] = alloc, cached
    gb = 1024**3
    print(f"RANK={rank} MEMSTATS {msg} device={torch.cuda.current_device()} "
          f"current alloc={alloc / gb:0.4f}GB (delta={d_alloc / gb:0.4f}GB max={torch.cuda.max_memory_allocated() / gb:0.4f}GB) "
          f"current cache={cached / gb:0.4f}GB (delta={d_cached / gb:0.4f}GB max={torch.cuda.max_memory_reserved() / gb:0.4f}GB)")


def get_ma_status():
    if dist.is_initialized() and dist.get_rank() != 0:
        return 0
    return torch.cuda.memory_allocated() if torch.cuda.is_available() else 0


def empty_cache():
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
    gc.collect()


_graph_cache = {}


def graph_process(replay_first_step, func, *args, **kwargs):
    """Capture ``func(*args)`` (device-only work on fixed addresses) into a CUDA graph on first use, replay afterwards."""
    key = func.__name__
    if key not in _graph_cache:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            func(*args, **kwargs)
        torch.cuda.current_stream().wait_stream(side)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            func(*args, **kwargs)
        _graph_cache[key] = g
        if replay_first_step:
            g.replay()
    else:
        _graph_cache[key].replay()


def _folder_size(folder):
    return sum(os.path.getsize(os.path.join(r, f)) for r, _, fs in os.walk(folder) for f in fs)


def get_checkpoint_folder_size(save_dir, tag, local_rank=None):
    """Bytes of ``<save_dir>/<tag>`` summed over nodes (local rank 0 of every node measures its folder)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
    size = _folder_size(os.path.join(save_dir, str(tag))) if local_rank == 0 else 0
    t = torch.tensor(size, dtype=torch.int64, device=dev)
    if dist.is_initialized():
        dist.reduce(t, dst=0)
    return int(t)


class TLinear(torch.nn.Linear):
    """Linear built from a layer whose weight is stored transposed ([in, out], e.g. GPT-2 Conv1D)."""

    def __init__(self, orig_layer, name=""):
        self.name = name
        super().__init__(orig_layer.weight.shape[1], orig_layer.weight.shape[0], bias=orig_layer.bias is not None)
        self.weight.data = orig_layer.weight.data.t().contiguous()
        self.bias = orig_layer.bias


def _offload_key(key):
    return f"{key}_offload_buffer"


def offload_adam_states(optimizer, device, pin_memory=False, non_blocking=False):
    """Move ``exp_avg`` / ``exp_avg_sq`` of an Adam-style optimizer to ``device`` through persistent (optionally
    pinned) buffers kept in the state dict."""
    for state in optimizer.state.values():
        for key in ("exp_avg", "exp_avg_sq"):
            if key not in state:
                continue
            bk = _offload_key(key)
            if bk not in state:
                state[bk] = torch.empty_like(state[key], device=device)
                if pin_memory and torch.cuda.is_available():
                    state[bk] = state[bk].pin_memory()
            state[bk].copy_(state[key], non_blocking=non_blocking)
            state[key].data = state[bk]


def reload_adam_states(optimizer, device, non_blocking=False):
    for state in optimizer.state.values():
        for key in ("exp_avg", "exp_avg_sq"):
            if key in state and _offload_key(key) in state:
                state[key].data = state[_offload_key(key)].to(device, non_blocking=non_blocking)


def is_transformers_cache(obj):
    try:
        from transformers.cache_utils import Cache
    except ImportError:
        return False
    return isinstance(obj, Cache)


def compare_tensors_in_structures(inputs1, inputs2):
    """Deep equality of two lists / dicts that may contain tensors (HF cache objects compare by type)."""
    if type(inputs1) is not type(inputs2):
        return False

    def same(a, b):
        if is_transformers_cache(a) and is_transformers_cache(b):
            return type(a) is type(b)
        if torch.is_tensor(a) and torch.is_tensor(b):
            return a.shape == b.shape and torch.equal(a.to(b.device), b)
        if isinstance(a, (list, dict)) and isinstance(b, (list, dict)):
            return compare_tensors_in_structures(a, b)
        return a == b

    if isinstance(inputs1, list):
        return len(inputs1) == len(inputs2) and all(same(a, b) for a, b in zip(inputs1, inputs2))
    if isinstance(inputs1, dict):
        return inputs1.keys() == inputs2.keys() and all(same(inputs1[k], inputs2[k]) for k in inputs1)
    return False


def maybe_loss_for_backward(value):
    """A scalar tensor attached to a graph."""
    return torch.is_tensor(value) and value.numel() == 1 and value.grad_fn is not None


class OutputBackwardHookManager:
    """Runs ``preprocess_once_fn()`` when the FIRST gradient of a forward's outputs arrives (start of backward) and
    ``preprocess_per_tensor_fn(grad)`` for every output gradient. Lives as long as the outputs it is attached to."""

    def __init__(self, preprocess_once_fn, preprocess_per_tensor_fn=None):
        self.once, self.per_tensor = preprocess_once_fn, preprocess_per_tensor_fn
        self._fired = False
        self.handles = []

    def _hook(self, grad):
        if not self._fired:
            self._fired = True
            self.once()
        if self.per_tensor is not None:
            out = self.per_tensor(grad)
            return grad if out is None else out
        return grad

    def reset(self):
        self._fired = False

    def register_hooks_on_outputs(self, outputs):
        def visit(t):
            if t.requires_grad:
                self.handles.append(t.register_hook(self._hook))
                t._ds_backward_hook_manager = self  # ties the manager's lifetime to the tensor
            return t

        _map_structure(outputs, visit)
        return outputs

    def remove_hooks(self):
        for h in self.handles:
            h.remove()
        self.handles = []


def register_output_backward_hooks(outputs, preprocess_once_fn, preprocess_per_tensor_fn=None):
    mgr = OutputBackwardHookManager(preprocess_once_fn, preprocess_per_tensor_fn)
    mgr.register_hooks_on_outputs(outputs)
    return mgr


def check_internal_apis_for_count_used_parameters():
    """The counting below needs two private torch entry points."""
    try:
        from torch.autograd.graph import _get_grad_fn_or_grad_acc  # noqa: F401
    except ImportError:
        return False
    return hasattr(torch._C, "_current_graph_task_id") and hasattr(torch._C, "_will_engine_execute_node")


def count_used_parameters_in_backward(parameters):
    """How many of ``parameters`` will receive a gradient in the backward pass that is running right now (call it from
    inside a backward hook). ZeRO uses it to know how many gradient hooks to wait for."""
    assert check_internal_apis_for_count_used_parameters(), (
        "count_used_parameters_in_backward requires internal PyTorch APIs that are not available in this PyTorch build.")
    from torch.autograd.graph import _get_grad_fn_or_grad_acc
    if torch._C._current_graph_task_id() == -1:
        raise RuntimeError("count_used_parameters_in_backward must be called during backward execution")
    seen = set()
    for p in parameters:
        if not torch.is_tensor(p) or not p.requires_grad:
            continue
        node = _get_grad_fn_or_grad_acc(p)
        if node is not None:
            seen.add(node)
    return sum(1 for n in seen if torch._C._will_engine_execute_node(n))
