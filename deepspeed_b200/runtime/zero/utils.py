"""ZeRO helper functions (reference ``runtime/zero/utils.py``)."""
import gc
from typing import List, Tuple

import torch

from ... import comm as dist
from ...utils.logging import logger


class ZeRORuntimeException(Exception):
    pass


def _supported_optimizers():
    from ...ops.adam import DeepSpeedCPUAdam, FusedAdam, ZenFlowCPUAdam
    from ...ops.cpu_optimizers import DeepSpeedCPUAdagrad, DeepSpeedCPULion
    from ...ops.optimizers import FusedLion
    out = [torch.optim.Adam, torch.optim.AdamW, FusedAdam, DeepSpeedCPUAdam, ZenFlowCPUAdam, torch.optim.Adagrad,
           DeepSpeedCPUAdagrad, DeepSpeedCPULion, FusedLion]
    try:
        from ...ops.muon import Muon
        out.append(Muon)
    except Exception:  # noqa: BLE001
        pass
    return out


def __getattr__(name):
    if name == "ZERO_SUPPORTED_OPTIMIZERS":
        return _supported_optimizers()
    raise AttributeError(name)


def is_zero_supported_optimizer(optimizer):
    """Optimizers whose update is element-wise, so that sharding their state is exact (anything else needs
    ``zero_allow_untested_optimizer``)."""
    if not dist.is_initialized() or dist.get_rank() == 0:
        logger.info(f"Checking ZeRO support for optimizer={optimizer.__class__.__name__} type={type(optimizer)}")
    return type(optimizer) in _supported_optimizers()


def _device():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def get_lst_from_rank0(lst: List[int]) -> List[int]:
    """Rank 0's list on every rank (e.g. the trace of parameter ids, to detect ranks that diverged)."""
    t = torch.tensor(lst if dist.get_rank() == 0 else [-1] * len(lst), dtype=torch.int, device=_device())
    dist.broadcast(t, src=0)
    return t.cpu().tolist()


def _fingerprints(lst):
    """(length, order-sensitive checksum) of every rank's list, gathered on all ranks in ONE collective: every rank
    reaches the same verdict, so a disagreement raises everywhere instead of leaving the agreeing ranks stuck in the
    next collective."""
    h = 0
    for i, v in enumerate(lst):
        h = (h * 1000003 + int(v) + i) % 2147483629
    mine = torch.tensor([len(lst), h], dtype=torch.int64, device=_device())
    out = [torch.zeros_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [tuple(int(x) for x in t.tolist()) for t in out]


def assert_lst_len_same_as_other_ranks(lst: List[int]) -> None:
    fp = _fingerprints(lst)
    bad = [r for r, f in enumerate(fp) if f[0] != fp[0][0]]
    if bad:
        raise RuntimeError(f"disagreement between rank0 and rank{bad[0]}: rank0: {fp[0][0]}, rank{bad[0]}: "
                           f"{fp[bad[0]][0]} (list lengths)")


def assert_ints_same_as_other_ranks(ints: List[int]) -> None:
    """Every rank must hold the same list as rank 0 (otherwise collectives would be issued in different orders)."""
    fp = _fingerprints(ints)
    bad = [r for r, f in enumerate(fp) if f != fp[0]]
    if bad:
        r = bad[0]
        what = "lengths" if fp[r][0] != fp[0][0] else "contents"
        mine = list(ints) if len(ints) <= 32 else list(ints[:32]) + ["..."]
        raise RuntimeError(f"disagreement between rank0 and rank{r} ({what}): rank0 (len, checksum) = {fp[0]}, "
                           f"rank{r} = {fp[r]}; this rank ({dist.get_rank()}) holds {mine}")


def is_builtin_type(obj):
    return obj.__class__.__module__ in ("__builtin__", "builtins")


def isinstance_namedtuple(obj) -> bool:
    return isinstance(obj, tuple) and hasattr(obj, "_asdict") and hasattr(obj, "_fields")


def is_zero_param(parameter):
    if not torch.is_tensor(parameter):
        return False
    return hasattr(parameter, "ds_id") or hasattr(parameter, "ds_unit")


_warned = [False]


def apply_to_tensors_only(function, value, warning_msg_fn=None):
    """Apply ``function`` to every tensor inside a nested list / tuple / namedtuple / dict; other leaves pass through
    (with one warning for non-builtin types when ``warning_msg_fn`` is given)."""
    if isinstance(value, (tuple, list)):
        out = [apply_to_tensors_only(function, v) for v in value]
        return value.__class__(*out) if isinstance_namedtuple(value) else value.__class__(out)
    if isinstance(value, dict):
        for k in value:
            value[k] = apply_to_tensors_only(function, value[k])
        return value
    if torch.is_tensor(value):
        touched = function(value)
        if is_zero_param(value) and not is_zero_param(touched):
            touched.ds_param_alias = value
        return touched
    if not is_builtin_type(value) and warning_msg_fn and not _warned[0] and (not dist.is_initialized()
                                                                             or dist.get_rank() == 0):
        logger.warning(warning_msg_fn(value))
        _warned[0] = True
    return value


def get_mapping_to_flat_buffer(tensors: List[torch.Tensor]) -> List[Tuple[torch.Tensor, int, int]]:
    out, off = [], 0
    for t in tensors:
        out.append((t, off, t.numel()))
        off += t.numel()
    return out


def defragment(tensors: List[torch.Tensor]) -> torch.Tensor:
    """Move the tensors into ONE contiguous device buffer (bouncing through host memory so the old fragments are
    released before the buffer is allocated); the tensors become views of it."""
    assert len({t.dtype for t in tensors}) == 1 and len({t.device for t in tensors}) == 1
    device = tensors[0].device
    host = torch.empty(sum(t.numel() for t in tensors), dtype=tensors[0].dtype, device="cpu")
    infos = get_mapping_to_flat_buffer(tensors)
    for t, off, n in infos:
        host.narrow(0, off, n).copy_(t.reshape(-1))
        t.data = torch.empty(0, dtype=t.dtype, device=t.device)
    gc.collect()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
    flat = host.to(device)
    for t, off, n in infos:
        t.data = flat.narrow(0, off, n)
    return flat


# ---- small helpers of the reference's stage_1_and_2.py / stage3.py / partition_parameters.py -----------------------
def split_half_float_double(tensors):
    """Tensors bucketed by dtype (fp16, fp32, fp64, bf16 order): one flat collective per bucket."""
    order = (torch.float16, torch.float32, torch.float64, torch.bfloat16)
    buckets = [[t for t in tensors if t.dtype == dt] for dt in order]
    return [b for b in buckets if b]


def split_half_float_double_sparse(tensors):
    """Like :func:`split_half_float_double`, with sparse gradients (``SparseTensor``) in their own buckets first.
    Returns ``[(dtype, [tensors]), ...]`` (reference ``runtime/engine.py``)."""
    from ..sparse_tensor import SparseTensor
    sparse = [t for t in tensors if isinstance(t, SparseTensor)]
    dense = [t for t in tensors if not isinstance(t, SparseTensor)]
    out = []
    if sparse:
        out.append((SparseTensor.type(), sparse))
    for dt in (torch.float16, torch.float32, torch.float64, torch.bfloat16):
        b = [t for t in dense if t.dtype == dt]
        if b:
            out.append((dt, b))
    return out


def isclose(a, b, rtol=1e-09, atol=0.0):
    return abs(a - b) <= max(rtol * max(abs(a), abs(b)), atol)


def lcm(x, y):
    from math import gcd
    return x * y // gcd(x, y)


def get_alignment_padding(tensor_list, alignment):
    """Elements to append so that the tensors' total count is a multiple of ``alignment``."""
    rem = sum(t.numel() for t in tensor_list) % alignment
    return (alignment - rem) if rem else 0


def move_to_cpu(tensor_list):
    for t in tensor_list:
        t.data = t.data.cpu()


def print_rank_msg(msg):
    from ... import comm as dist
    print(f"rank {dist.get_rank() if dist.is_initialized() else 0} - {msg}")


def get_allgather_dtype(param, param_ds_tensor):
    """Dtype a parameter travels in: its autocast dtype when torch autocast manages it, else that of its shard."""
    dt = getattr(param, "autocast_dtype", None) or getattr(param, "ds_autocast_dtype", None)
    return dt if dt is not None else param_ds_tensor.dtype


def get_new_tensor_fn_for_dtype(dtype):
    """``Tensor.new_tensor`` that casts floating-point results to ``dtype`` (used while ``zero.Init`` is active)."""

    def new_tensor(cls, *args, **kwargs):
        t = torch.Tensor.new_tensor(cls, *args, **kwargs)
        return t.to(dtype) if t.is_floating_point() else t

    return new_tensor


def zero_wrapper_for_fp_tensor_constructor(fn, target_fp_dtype):
    """Wrap a tensor constructor (``torch.empty`` ...) so that floating-point tensors come out in ``target_fp_dtype``
    on the current accelerator unless a device was given."""

    def wrapped(*args, **kwargs):
        if kwargs.get("device") is None and torch.cuda.is_available():
            kwargs["device"] = torch.device("cuda", torch.cuda.current_device())
        t = fn(*args, **kwargs)
        return t.to(target_fp_dtype) if t.is_floating_point() else t

    return wrapped
