"""``PartitionedOptimizerSwapper``: synchronous swap-in before / swap-out after each sub-group's optimizer step
(reference ``runtime/swap_tensor/partitioned_optimizer_swapper.py:26``)."""
import torch

from .optimizer_utils import OptimizerSwapper
from .utils import swap_in_tensors, swap_out_tensors


class PartitionedOptimizerSwapper(OptimizerSwapper):

    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers=None):
        super().__init__(swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers)
        count = max(1, int(getattr(swap_config, "buffer_count", 4)))
        self.buffers = [self.new_pinned(self.largest_numel) for _ in range(count)]

    def initialize_parameters(self, parameters, src_tensors):
        """Write the initial fp32 values of ``parameters`` to their swap files and release the memory."""
        for p, src in zip(parameters, src_tensors):
            info = self._get_param_swap_info(p) or self._create_param_swap_info(p, src.numel())
            buf = self.buffers[0].narrow(0, 0, self._io_aligned_numel(src.numel()))
            buf.narrow(0, 0, src.numel()).copy_(src.view(-1))
            swap_out_tensors(self.aio_handle, [buf], [info.swap_paths[0]])
            self.aio_handle.wait()
            p.data = torch.Tensor()

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        info = self._get_param_swap_info(parameter)
        if info is None:
            return
        self._register_states(info, parameter)
        need = len(info.tensors)
        assert need <= len(self.buffers), f"sub-group needs {need} swap buffers, {len(self.buffers)} configured"
        n = self._io_aligned_numel(info.numel())
        bufs = [b.narrow(0, 0, n) for b in self.buffers[:need]]
        swap_in_tensors(self.aio_handle, bufs, info.swap_paths[:need])
        self.aio_handle.wait()
        info.set_swap_buffers(bufs)

    def swap_out_optimizer_state(self, parameter, async_swap=False):
        info = self._get_param_swap_info(parameter)
        if info is None:
            return
        self._register_states(info, parameter)
        n = self._io_aligned_numel(info.numel())
        bufs = []
        for i, t in enumerate(info.tensors):
            b = self.buffers[i].narrow(0, 0, n)
            if t.data_ptr() != b.data_ptr():
                b.narrow(0, 0, info.numel()).copy_(t.view(-1))
            bufs.append(b)
        swap_out_tensors(self.aio_handle, bufs, info.swap_paths)
        self.aio_handle.wait()
        info.release_memory()

    def swap_out_gradients(self, parameter, gradient_offsets, gradient_tensors):
        info = self._get_param_swap_info(parameter)
        if info is None:
            return
        paths = info.get_or_create_gradient_paths(gradient_offsets, [g.numel() for g in gradient_tensors])
        for g, path in zip(gradient_tensors, paths):
            n = self._io_aligned_numel(g.numel())
            buf = self.buffers[-1].narrow(0, 0, n)
            buf.narrow(0, 0, g.numel()).copy_(g.view(-1))
            swap_out_tensors(self.aio_handle, [buf], [path])
            self.aio_handle.wait()
