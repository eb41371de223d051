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


def assert_lst_len_same_as_other_ranks(lst: List[int]) -> None:
    got = get_lst_from_rank0([len(lst)])[0]
    if got != len(lst):
        raise RuntimeError(f"disagreement between rank0 and rank{dist.get_rank()}: rank0: {got}, "
                           f"rank{dist.get_rank()}: {len(lst)}")


def assert_ints_same_as_other_ranks(ints: List[int]) -> None:
    """Every rank must hold the same list as rank 0 (otherwise collectives would be issued in different orders)."""
    assert_lst_len_same_as_other_ranks(ints)
    r0 = get_lst_from_rank0(ints)
    if r0 != list(ints):
        raise RuntimeError(f"disagreement between rank0 and rank{dist.get_rank()}: rank0: {r0}, "
                           f"rank{dist.get_rank()}: {ints}")


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
