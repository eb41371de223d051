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


def can_send_recv() -> bool:
    """Point-to-point send / recv is what the transport uses (no 2-rank-broadcast fallback needed)."""
    return True


def send_obj(msg, dest: int):
    """Send an arbitrary msgpack-serialisable Python object to global rank ``dest`` (length first, then the bytes).
    Reference ``runtime/pipe/p2p.py:96``. The payload travels through a device tensor: use sparingly."""
    import msgpack
    raw = msgpack.packb(msg)
    payload = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(_dev())
    dist.send(torch.tensor([payload.numel()], dtype=torch.long, device=_dev()), dest)
    dist.send(payload, dest)


def recv_obj(sender: int):
    """Counterpart of :func:`send_obj`; tensors found in the object are moved to the current device."""
    import msgpack
    length = torch.zeros(1, dtype=torch.long, device=_dev())
    dist.recv(length, sender)
    payload = torch.empty(int(length.item()), dtype=torch.uint8, device=_dev())
    dist.recv(payload, sender)
    msg = msgpack.unpackb(payload.cpu().numpy().tobytes())

    def to_device(x):
        if torch.is_tensor(x):
            return x.to(_dev())
        if isinstance(x, (list, tuple)):
            return type(x)(to_device(v) for v in x)
        if isinstance(x, dict):
            return {k: to_device(v) for k, v in x.items()}
        return x

    return to_device(msg)


def wait():
    """Block until the p2p operations issued so far have completed."""
    if torch.cuda.is_available():
        torch.cuda.current_stream().synchronize()
