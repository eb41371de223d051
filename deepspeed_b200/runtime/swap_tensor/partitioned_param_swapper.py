"""``AsyncPartitionedParameterSwapper``: ZeRO-3 parameter shards on NVMe (reference
``partitioned_param_swapper.py:37``). Each parameter (unit shard) has a swap file; a bounded pool of pinned buffers
holds the shards that are currently "available"; reads are asynchronous so the coordinator can prefetch."""
import os
from enum import Enum

import torch

from ...ops.aio import aio_handle as AioHandle
from .utils import AIO_ALIGNED_BYTES, MIN_AIO_BYTES, pinned


class PartitionedParamStatus(Enum):
    AVAILABLE = 1      # resident in a pinned buffer
    NOT_AVAILABLE = 2  # only on NVMe
    INFLIGHT = 3       # read issued


class AsyncPartitionedParameterSwapper:

    def __init__(self, ds_config, model_dtype):
        from .aio_config import get_aio_config
        zc = ds_config.zero_config if hasattr(ds_config, "zero_config") else ds_config
        op = zc.offload_param
        aio = get_aio_config(getattr(ds_config, "_param_dict", {}))
        self.dtype = model_dtype
        rank = int(os.environ.get("RANK", "0"))
        self.swap_folder = os.path.join(str(op.nvme_path), "zero_stage_3", f"{str(model_dtype).split('.')[-1]}params",
                                        f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.swap_element_size = torch.tensor([], dtype=model_dtype).element_size()
        self.aligned_bytes = AIO_ALIGNED_BYTES * aio["intra_op_parallelism"]
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.min_aio_bytes = max(MIN_AIO_BYTES, aio["block_size"])
        self.elements_per_buffer = int(op.buffer_size)
        self.param_buffer_count = int(op.buffer_count)
        self.buffers = [pinned(self._aligned(self.elements_per_buffer), model_dtype)
                        for _ in range(self.param_buffer_count)]
        self.available_buffer_ids = list(range(self.param_buffer_count))
        self.id_to_path, self.param_id_to_buffer_id, self.param_id_to_numel = {}, {}, {}
        self.status = {}
        self.pending_reads, self.pending_writes = 0, 0
        self.inflight_params = []
        self.available_params = set()
        self.available_numel = 0
        self.aio_read_handle = AioHandle(aio["block_size"], aio["queue_depth"], aio["single_submit"],
                                         aio["overlap_events"], aio["intra_op_parallelism"])
        self.aio_write_handle = AioHandle(aio["block_size"], aio["queue_depth"], aio["single_submit"],
                                          aio["overlap_events"], aio["intra_op_parallelism"])
        self.invalid_buffer = torch.tensor(1).half()

    def _aligned(self, numel):
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    @staticmethod
    def _pid(param):
        return getattr(param, "ds_id", id(param))

    def available_swap_in_buffers(self):
        return len(self.available_buffer_ids)

    def swappable_tensor(self, param=None, numel=None):
        numel = numel if numel is not None else getattr(param, "ds_numel", param.numel())
        return self.min_aio_bytes <= numel * self.swap_element_size

    def get_path(self, param, must_exist=False):
        pid = self._pid(param)
        if pid not in self.id_to_path:
            assert not must_exist, f"no swap file for parameter {pid}"
            self.id_to_path[pid] = os.path.join(self.swap_folder, f"{pid}_param.tensor.swp")
        return self.id_to_path[pid]

    def get_buffer(self, param, numel):
        """Reserve a pinned buffer for ``param`` (e.g. to write its freshly initialised shard)."""
        pid = self._pid(param)
        assert self.available_buffer_ids, "no free parameter swap buffers"
        assert numel <= self.elements_per_buffer, f"shard of {numel} elements exceeds offload_param.buffer_size"
        bid = self.available_buffer_ids.pop()
        self.param_id_to_buffer_id[pid], self.param_id_to_numel[pid] = bid, numel
        self.status[pid] = PartitionedParamStatus.AVAILABLE
        return self.buffers[bid].narrow(0, 0, numel)

    def swap_out_and_release(self, params, async_op=False, force_buffer_release=False):
        for p in params:
            pid = self._pid(p)
            bid = self.param_id_to_buffer_id[pid]
            n = self._aligned(self.param_id_to_numel[pid])
            assert self.aio_write_handle.async_pwrite(self.buffers[bid].narrow(0, 0, n), self.get_path(p)) == 0
            self.pending_writes += 1
        if not async_op or force_buffer_release:
            self.synchronize_writes()
        for p in params:
            self.remove_partition_and_release_buffers([p])

    def swap_in(self, params, async_op=True, swap_in_buffers=None):
        for p in params:
            pid = self._pid(p)
            if self.status.get(pid) in (PartitionedParamStatus.AVAILABLE, PartitionedParamStatus.INFLIGHT):
                continue
            assert self.available_buffer_ids, "no free parameter swap buffers for swap_in"
            bid = self.available_buffer_ids.pop()
            self.param_id_to_buffer_id[pid] = bid
            n = self._aligned(self.param_id_to_numel[pid])
            assert self.aio_read_handle.async_pread(self.buffers[bid].narrow(0, 0, n), self.get_path(p, True)) == 0
            self.status[pid] = PartitionedParamStatus.INFLIGHT
            self.inflight_params.append(p)
            self.pending_reads += 1
        if not async_op:
            self.synchronize_reads()

    def synchronize_reads(self):
        if self.pending_reads:
            self.aio_read_handle.wait()
            self.pending_reads = 0
        for p in self.inflight_params:
            pid = self._pid(p)
            self.status[pid] = PartitionedParamStatus.AVAILABLE
            self.available_params.add(pid)
        self.inflight_params = []

    def synchronize_writes(self):
        if self.pending_writes:
            self.aio_write_handle.wait()
            self.pending_writes = 0

    def tensor(self, param):
        """The resident shard of ``param`` as a host tensor view."""
        pid = self._pid(param)
        if self.status.get(pid) == PartitionedParamStatus.INFLIGHT:
            self.synchronize_reads()
        assert self.status.get(pid) == PartitionedParamStatus.AVAILABLE
        return self.buffers[self.param_id_to_buffer_id[pid]].narrow(0, 0, self.param_id_to_numel[pid])

    def remove_partition_and_release_buffers(self, params):
        for p in params:
            pid = self._pid(p)
            bid = self.param_id_to_buffer_id.pop(pid, None)
            if bid is not None:
                self.available_buffer_ids.append(bid)
            self.available_params.discard(pid)
            self.status[pid] = PartitionedParamStatus.NOT_AVAILABLE

    def release_and_reset_all(self, module=None):
        self.synchronize_reads()
        self.synchronize_writes()
        for pid in list(self.param_id_to_buffer_id):
            self.available_buffer_ids.append(self.param_id_to_buffer_id.pop(pid))
            self.status[pid] = PartitionedParamStatus.NOT_AVAILABLE
        self.available_params.clear()
