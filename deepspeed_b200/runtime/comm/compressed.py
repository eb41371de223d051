"""Error-compensated 1-bit all-reduce (reference ``runtime/comm/nccl.py:52`` ``NcclBackend.compressed_allreduce``,
``runtime/comm/compressed.py``; the reference packs bits with cupy).

Two compression stages, each with error feedback: every worker sends ``sign(x + e_w)`` (1 bit/element) plus one scale;
rank r, server for chunk r, averages the workers' reconstructions, compresses the average again (server error ``e_s``)
and all-gathers the result. Bits are packed 8 per byte with integer tensor ops on the device the data lives on."""
import numpy as np
import torch

from ... import comm as dist

_POW2 = {}


def _pow2(device):
    t = _POW2.get(device)
    if t is None:
        t = _POW2[device] = torch.tensor([1, 2, 4, 8, 16, 32, 64, 128], dtype=torch.uint8, device=device)
    return t


def pack_signs(x):
    """bit i of byte j = (x[8 j + i] >= 0). ``x.numel()`` must be a multiple of 8."""
    bits = (x >= 0).view(-1, 8).to(torch.uint8)
    return (bits * _pow2(x.device)).sum(dim=1, dtype=torch.uint8)


def unpack_signs(packed, dtype=torch.float32):
    """-> tensor of +-1 with ``8 * packed.numel()`` elements."""
    bits = (packed.view(-1, 1) & _pow2(packed.device)).ne(0)
    return bits.to(dtype).mul_(2).sub_(1).view(-1)


class CompressedBackend:

    def __init__(self, mpu=None):
        self.world_group = mpu.get_data_parallel_group() if mpu is not None else None
        self.rank = dist.get_rank(group=self.world_group)
        self.size = dist.get_world_size(group=self.world_group)

    def padded_size(self, numel):
        """Element count the error buffers need: a multiple of 8 bits per server chunk."""
        div = self.size * int(self.size * 8 / np.gcd(self.size, 8))  # same rounding as the reference
        return (numel + div - 1) // div * div

    def server_chunk(self, numel):
        return self.padded_size(numel) // self.size

    @torch.no_grad()
    def compressed_allreduce(self, buffer_m, worker_error, server_error, local_rank=None):
        """Returns the (lossy) average of ``buffer_m`` over the group, in ``buffer_m``'s shape; updates both errors."""
        shape, n = buffer_m.shape, buffer_m.numel()
        padded = worker_error.numel()
        flat = buffer_m.reshape(-1).float()
        if padded != n:
            flat = torch.cat([flat, flat.new_zeros(padded - n)])
        else:
            flat = flat.clone()
        # ---- worker compression ----
        flat.add_(worker_error)
        worker_scale = flat.norm() / np.sqrt(flat.numel())
        signs = torch.where(flat >= 0, 1.0, -1.0)
        worker_error.copy_(flat - worker_scale * signs)
        packed = pack_signs(flat).view(self.size, -1)
        recv = torch.empty_like(packed)
        scales = [torch.zeros_like(worker_scale) for _ in range(self.size)]
        if self.size > 1:
            dist.all_to_all_single(recv, packed.contiguous(), group=self.world_group)
            dist.all_gather(scales, worker_scale, group=self.world_group)
        else:
            recv, scales = packed, [worker_scale]
        # ---- server: average my chunk, compress again ----
        chunk = unpack_signs(recv).view(self.size, -1)
        comp = (chunk * torch.stack(scales).view(-1, 1)).sum(0).div_(self.size)
        comp.add_(server_error)
        server_scale = comp.norm() / np.sqrt(comp.numel())
        s_signs = torch.where(comp >= 0, 1.0, -1.0)
        server_error.copy_(comp - server_scale * s_signs)
        s_packed = pack_signs(comp)
        all_packed = torch.empty(self.size * s_packed.numel(), dtype=torch.uint8, device=flat.device)
        s_scales = [torch.zeros_like(server_scale) for _ in range(self.size)]
        if self.size > 1:
            dist.all_gather_into_tensor(all_packed, s_packed, group=self.world_group)
            dist.all_gather(s_scales, server_scale, group=self.world_group)
        else:
            all_packed, s_scales = s_packed, [server_scale]
        out = unpack_signs(all_packed).view(self.size, -1) * torch.stack(s_scales).view(-1, 1)
        return out.view(-1)[:n].view(shape).to(buffer_m.dtype)


NcclBackend = CompressedBackend
