"""Reduce-scatter of a list of tensors in one collective (reference ``runtime/comm/coalesced_collectives.py:158``) and
the qgZ quantized all-to-all reduction (:31 ``all_to_all_quant_reduce``)."""
import math

import torch

from ... import comm as dist
from ...ops.quantizer import all_to_all_loco_quant_reduce, all_to_all_quant_reduce  # noqa: F401


@torch.no_grad()
def reduce_scatter_coalesced(tensors, group=None):
    """Every tensor is split evenly (zero padded) over the group; rank r gets the sum of everyone's r-th pieces, for
    all tensors, with one reduce-scatter. Returns the list of this rank's partitions (averaged)."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    parts = [math.ceil(t.numel() / world) for t in tensors]
    dtype, device = tensors[0].dtype, tensors[0].device
    total = sum(parts)
    send = torch.zeros(world, total, dtype=dtype, device=device)
    off = 0
    for t, p in zip(tensors, parts):
        flat = t.reshape(-1)
        padded = torch.zeros(world * p, dtype=dtype, device=device)
        padded[:flat.numel()] = flat
        send[:, off:off + p] = padded.view(world, p)
        off += p
    send.div_(world)
    recv = torch.empty(total, dtype=dtype, device=device)
    if world > 1:
        dist.reduce_scatter_tensor(recv, send.view(-1), group=group)
    else:
        recv.copy_(send.view(-1))
    out, off = [], 0
    for t, p in zip(tensors, parts):
        lo = rank * p
        valid = max(0, min(p, t.numel() - lo))
        out.append(recv[off:off + valid])
        off += p
    return out
