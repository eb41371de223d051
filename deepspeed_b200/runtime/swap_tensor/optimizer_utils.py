"""Bookkeeping shared by the optimizer-state swappers (reference ``runtime/swap_tensor/optimizer_utils.py``): which
tensors belong to a parameter (its fp32 value + every optimizer state), their swap files, and the aligned sizes."""
import os

import torch

from ...ops.aio import aio_handle as AioHandle
from .utils import MIN_AIO_BYTES, AIO_ALIGNED_BYTES, pinned, swap_file


class FlattenedTensorSwapInfo:

    def __init__(self, path, length, offset):
        self.path, self.length, self.offset = path, length, offset


class OptimizerStateSwapInfo:
    """All swappable tensors of one parameter (sub-group): index 0 is the fp32 parameter, the rest optimizer states."""

    def __init__(self, parameter, numel, base_folder):
        self.tensors = []
        self.param_id = OptimizerSwapper.parameter_id(parameter)
        self.swap_folder = base_folder
        self.swap_paths = []
        self.swapped_gradients = {}
        self.unswapped_gradients = {}
        self.tensor_numel = numel
        self.tensor_dtype = parameter.dtype
        self.tensor_device = parameter.device
        self.has_state_tensors = False
        self._add_tensors([parameter])

    def numel(self):
        return self.tensor_numel

    def has_gradients(self):
        return bool(self.swapped_gradients) or bool(self.unswapped_gradients)

    def _add_tensors(self, tensor_list):
        for t in tensor_list:
            self.tensors.append(t)
            self.swap_paths.append(os.path.join(self.swap_folder, f"{self.param_id}_{len(self.tensors) - 1}.tensor.swp"))

    def add_state_tensors(self, tensor_list):
        self.has_state_tensors = True
        self._add_tensors(tensor_list)

    def device(self):
        return self.tensor_device

    def dtype(self):
        return self.tensor_dtype

    def release_memory(self):
        for t in self.tensors:
            t.data = torch.Tensor()

    def get_or_create_gradient_paths(self, offsets, lengths):
        paths = []
        for off, ln in zip(offsets, lengths):
            if off not in self.swapped_gradients:
                path = os.path.join(self.swap_folder, f"{self.param_id}_gradient_{off}_{ln}.tensor.swp")
                self.swapped_gradients[off] = FlattenedTensorSwapInfo(path, ln, off)
            paths.append(self.swapped_gradients[off].path)
        return paths

    def set_swap_buffers(self, buffers):
        for t, b in zip(self.tensors, buffers):
            t.data = b.narrow(0, 0, self.numel()).data

    def get_swap_gradient_buffers(self, swap_buffer):
        return [swap_buffer.narrow(0, g.offset, g.length) for g in self.swapped_gradients.values()]

    def get_swap_gradient_paths(self):
        return [g.path for g in self.swapped_gradients.values()]

    def get_unpinned_state_tensors(self):
        return [t for t in self.tensors if t.numel() and not t.is_pinned()]

    def release_unswapped_gradients(self):
        self.unswapped_gradients = {}


class OptimizerSwapper:
    """Base: aligned sizes, swap folder ``<nvme_path>/zero_stage_3/optimizer/rank<r>``, the aio handle, and the
    param -> :class:`OptimizerStateSwapInfo` table."""

    @staticmethod
    def parameter_id(param):
        return getattr(param, "ds_id", id(param))

    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers=None):
        self.swap_config, self.aio_config = swap_config, aio_config
        self.optimizer, self.dtype, self.device, self.timers = optimizer, dtype, device, timers
        rank = int(os.environ.get("RANK", "0"))
        self.swap_folder = os.path.join(str(base_folder), "zero_stage_3", "optimizer", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio_config["block_size"])
        self.aligned_bytes = AIO_ALIGNED_BYTES * aio_config["intra_op_parallelism"]
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.largest_numel = self._io_aligned_numel(largest_numel)
        self.swap_params_info = {}
        self.aio_handle = AioHandle(aio_config["block_size"], aio_config["queue_depth"], aio_config["single_submit"],
                                    aio_config["overlap_events"], aio_config["intra_op_parallelism"])

    def _io_aligned_numel(self, numel):
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def swappable_tensor(self, param=None, numel=None):
        numel = param.numel() if numel is None else numel
        return self.min_aio_bytes <= numel * self.swap_element_size

    def _get_param_swap_info(self, parameter):
        return self.swap_params_info.get(self.parameter_id(parameter))

    def _create_param_swap_info(self, parameter, numel):
        info = OptimizerStateSwapInfo(parameter, numel, self.swap_folder)
        self.swap_params_info[self.parameter_id(parameter)] = info
        self._register_states(info, parameter)
        return info

    def _register_states(self, info, parameter):
        if not info.has_state_tensors:
            state = self.optimizer.state.get(parameter, {})
            tensors = [v for v in state.values() if torch.is_tensor(v) and v.numel() == info.numel()]
            if tensors:
                info.add_state_tensors(tensors)

    def new_pinned(self, numel):
        return pinned(self._io_aligned_numel(numel), self.dtype)

    def purge_state(self):
        for info in self.swap_params_info.values():
            info.tensors = [info.tensors[0]]
            info.has_state_tensors = False

    def swap_file(self, name):
        return swap_file(self.swap_folder, name)
