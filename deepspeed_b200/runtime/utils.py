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
