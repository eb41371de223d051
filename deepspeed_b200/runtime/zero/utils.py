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
