"""Pipeline p2p: send/recv of activation/grad tensors between adjacent stages over NCCL p2p.

Reference: ``runtime/pipe/p2p.py:46-94`` (``send/recv`` with the 2-rank-broadcast fallback) and the tensor-meta
handshake of ``pipe/engine.py:956-1045``. NCCL p2p stays the transport (north-star); transfers are issued on a side
stream so they overlap the neighbouring micro-batch's compute.
"""
import torch

from ... import comm as dist

_grid = None
_DTYPES = [torch.float32, torch.float16, torch.bfloat16, torch.int64, torch.int32, torch.bool, torch.uint8, torch.float64]


def init_process_groups(grid):
    global _grid
    _grid = grid


def _dev():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


def send(tensor, dest_stage, async_op=False):
    dst = _grid.stage_to_global(stage_id=dest_stage)
    if async_op:
        return dist.isend(tensor.contiguous(), dst)
    return dist.send(tensor.contiguous(), dst)


def recv(tensor, src_stage, async_op=False):
    src = _grid.stage_to_global(stage_id=src_stage)
    if async_op:
        return dist.irecv(tensor, src)
    return dist.recv(tensor, src)


def send_meta(obj, dest_stage):
    """Describe a tensor / tuple of tensors (count, ndims, dtype, shape) so the receiver can allocate buffers."""
    tensors = (obj, ) if torch.is_tensor(obj) else tuple(obj)
    header = [0 if torch.is_tensor(obj) else 1, len(tensors)]
    for t in tensors:
        header += [_DTYPES.index(t.dtype), t.dim()] + list(t.shape) + [int(t.requires_grad)]
    meta = torch.tensor([len(header)] + header, dtype=torch.int64, device=_dev())
    send(meta[:1].clone(), dest_stage)
    send(meta[1:].clone(), dest_stage)


def recv_meta(src_stage):
    n = torch.zeros(1, dtype=torch.int64, device=_dev())
    recv(n, src_stage)
    header = torch.zeros(int(n.item()), dtype=torch.int64, device=_dev())
    recv(header, src_stage)
    h = header.tolist()
    is_tuple, count = h[0], h[1]
    pos, specs = 2, []
    for _ in range(count):
        dt, nd = _DTYPES[h[pos]], h[pos + 1]
        shape = h[pos + 2:pos + 2 + nd]
        rg = bool(h[pos + 2 + nd])
        specs.append((dt, tuple(shape), rg))
        pos += 3 + nd
    return bool(is_tuple), specs
