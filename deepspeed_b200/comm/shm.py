"""Shared-memory all-reduce for CPU tensors when every rank is on this host.

Reference: ``comm/torch.py:186 inference_all_reduce`` -> ``torch.ops.deepspeed.inference_all_reduce_``
(``csrc/cpu/comm/shm_interface.cpp``, taken when ``LOCAL_SIZE == WORLD_SIZE``). Native side: ``csrc/cpu/shm_comm.cpp``.
"""
import atexit
import os

import torch
import torch.distributed as dist

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}
_STATE = {"handle": None, "tried": False}


def all_ranks_local():
    """LOCAL_SIZE (set by the launcher) equals the world size, or ``DSB200_SHM=1`` asserts it."""
    if os.environ.get("DSB200_SHM", "") == "0":
        return False
    if os.environ.get("DSB200_SHM", "") == "1":
        return True
    ls = os.environ.get("LOCAL_SIZE") or os.environ.get("LOCAL_WORLD_SIZE")
    return ls is not None and dist.is_initialized() and int(ls) == dist.get_world_size()


def _init(buf_bytes=32 << 20):
    """Collective over the world group: rank 0 creates the segment, the others attach after a barrier."""
    from ..ops import cpu_lib
    _STATE["tried"] = True
    L = cpu_lib.lib()
    if not hasattr(L, "dsb200_shm_create"):
        return None
    rank, world = dist.get_rank(), dist.get_world_size()
    name = f"/dsb200_shm_{os.environ.get('MASTER_PORT', '0')}_{world}".encode()
    h = L.dsb200_shm_create(name, rank, world, buf_bytes, 1, 120.0) if rank == 0 else None
    dist.barrier()
    if rank != 0:
        h = L.dsb200_shm_create(name, rank, world, buf_bytes, 0, 120.0)
    ok = torch.tensor([1 if h else 0], dtype=torch.int32)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if int(ok) == 0:  # someone could not map it (e.g. not the same host after all): nobody uses it
        if h:
            L.dsb200_shm_destroy(h)
        return None
    _STATE["handle"] = h
    atexit.register(shutdown)
    return h


def shutdown():
    h = _STATE.get("handle")
    if h:
        from ..ops import cpu_lib
        cpu_lib.lib().dsb200_shm_destroy(h)
        _STATE["handle"] = None


def shm_all_reduce(tensor):
    """In-place sum of a contiguous CPU tensor over all ranks. Returns False when the SHM path does not apply (the
    caller falls back to the backend all-reduce)."""
    if tensor.device.type != "cpu" or tensor.dtype not in _DT or not tensor.is_contiguous():
        return False
    if _STATE["handle"] is None:
        if _STATE["tried"] or not all_ranks_local():
            return False
        if _init() is None:
            return False
    from ..ops import cpu_lib
    rc = cpu_lib.lib().dsb200_shm_all_reduce(_STATE["handle"], tensor.data_ptr(), tensor.numel(), _DT[tensor.dtype])
    if rc != 0:
        raise RuntimeError(f"shm all-reduce failed ({rc}: {'peer timeout' if rc == -2 else 'bad argument'})")
    return True
