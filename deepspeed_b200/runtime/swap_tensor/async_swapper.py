"""``AsyncTensorSwapper``: fire-and-forget swap-out of tensors through a small pool of pinned staging buffers
(reference ``runtime/swap_tensor/async_swapper.py:18``). Used for gradient partitions during backward."""
import torch

from .utils import SwapBufferManager, swap_out_tensors

INVALID_BUFFER_INDEX = -1
ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"


class AsyncTensorSwapper:

    def __init__(self, aio_handle, numel_alignment, timers=None):
        self.aio_handle = aio_handle
        self.numel_alignment = numel_alignment
        self.timers = timers
        self.free_buffers, self.inflight = [], []  # staging tensors
        self.manager = None
        self.dtype = None
        self.num_pending_swaps = 0
        self.swapped_bytes = 0

    def has_buffers(self):
        return self.manager is not None

    def add_buffers(self, buffer_list):
        assert not self.has_buffers()
        self.dtype = buffer_list[0].dtype
        self.free_buffers = list(buffer_list)
        self.manager = True

    def get_timer_names(self):
        return [ASYNC_SWAPPER_WAIT_TIMER]

    def release_buffers(self):
        self._flush()
        out, self.free_buffers, self.manager = self.free_buffers, [], None
        return out

    def swap_out_tensors(self, tensor_list, path_list):
        for tensor, path in zip(tensor_list, path_list):
            self._swap_out_tensor(tensor, path)

    def _aligned(self, numel):
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def _swap_out_tensor(self, tensor, swap_path):
        assert self.has_buffers()
        n = self._aligned(tensor.numel())
        if not self.free_buffers:
            self._flush()
        staging = self.free_buffers.pop()
        assert n <= staging.numel(), "staging buffer smaller than the tensor"
        view = staging.narrow(0, 0, n)
        view.narrow(0, 0, tensor.numel()).copy_(tensor.detach().view(-1).to(self.dtype))
        swap_out_tensors(self.aio_handle, [view], [swap_path])
        self.inflight.append(staging)
        self.num_pending_swaps += 1
        self.swapped_bytes += n * view.element_size()

    def _flush(self):
        if self.num_pending_swaps:
            self.aio_handle.wait()
            self.num_pending_swaps = 0
        self.free_buffers.extend(self.inflight)
        self.inflight = []

    def wait(self):
        self._flush()


def make_staging(count, numel, dtype):
    mgr = SwapBufferManager(numel, count, dtype)
    return mgr.allocate_all(numel, dtype)
