"""GPUDirect Storage handle (reference ``csrc/gds`` over cuFile). cuFile (libcufile) is not part of this image, so the
GDS fast path is unavailable on this box; ``gds_handle`` therefore bounces through pinned host memory with the
libaio-based ``aio_handle`` (same API: device tensors are staged through a pinned buffer)."""
import torch

from ..aio import aio_handle


class gds_handle(aio_handle):

    def __init__(self, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True,
                 intra_op_parallelism=1):
        super().__init__(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism)
        self._bounce = {}

    def _stage(self, tensor):
        key = (tensor.numel(), tensor.dtype)
        if key not in self._bounce:
            self._bounce[key] = torch.empty(tensor.numel(), dtype=tensor.dtype, device="cpu",
                                            pin_memory=torch.cuda.is_available())
        return self._bounce[key]

    def pread(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        if buffer.device.type == "cpu":
            return super().pread(buffer, filename, validate, async_op, file_offset)
        host = self._stage(buffer)
        super().pread(host, filename, validate, False, file_offset)
        buffer.view(-1).copy_(host, non_blocking=async_op)
        return 0 if async_op else 1

    def pwrite(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        if buffer.device.type == "cpu":
            return super().pwrite(buffer, filename, validate, async_op, file_offset)
        host = self._stage(buffer)
        host.copy_(buffer.view(-1))
        return super().pwrite(host, filename, validate, async_op, file_offset)

    def new_pinned_device_tensor(self, num_elem, example_tensor):
        return torch.empty(num_elem, dtype=example_tensor.dtype, device=example_tensor.device)

    def free_pinned_device_tensor(self, tensor):
        return True

    def pin_device_tensor(self, tensor):
        return True

    def unpin_device_tensor(self, tensor):
        return True


class GDSBuilder:
    NAME = "gds"

    def is_compatible(self, verbose=False):
        return False  # cuFile is not installed; the handle above is the host-bounce fallback

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
