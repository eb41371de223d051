"""``PipelinedOptimizerSwapper``: double-buffered optimizer-state swapping, the read of sub-group i+1 and the write of
sub-group i-1 overlap the CPU optimizer step of sub-group i (reference ``pipelined_optimizer_swapper.py:52``).

The unit of swapping is a *state set*: a list of equally long fp32 tensors (moments, optionally the master weights)
identified by a key. Two aio handles (reads / writes) with their own thread pools keep both directions in flight."""
import os

import torch

from ...ops.aio import aio_handle as AioHandle
from .utils import AIO_ALIGNED_BYTES, pinned


class PipelinedOptimizerSwapper:

    def __init__(self, swap_config, aio_config, base_folder, largest_numel, tensors_per_set, dtype=torch.float32,
                 sets_in_flight=3):
        rank = int(os.environ.get("RANK", "0"))
        self.folder = os.path.join(str(base_folder), "zero_stage_3", "optimizer", f"rank{rank}")
        os.makedirs(self.folder, exist_ok=True)
        esz = torch.tensor([], dtype=dtype).element_size()
        self.align = max(1, AIO_ALIGNED_BYTES * aio_config["intra_op_parallelism"] // esz)
        self.numel = (largest_numel + self.align - 1) // self.align * self.align
        self.tensors_per_set, self.dtype = tensors_per_set, dtype
        mk = lambda: AioHandle(aio_config["block_size"], aio_config["queue_depth"], aio_config["single_submit"],  # noqa
                               aio_config["overlap_events"], aio_config["intra_op_parallelism"])
        self.read_handle, self.write_handle = mk(), mk()
        self.slots = [[pinned(self.numel, dtype) for _ in range(tensors_per_set)] for _ in range(sets_in_flight)]
        self.free_slots = list(range(sets_in_flight))
        self.loaded = {}      # key -> slot (read finished or in flight)
        self.reading = None   # key whose read is in flight
        self.writing = []     # slots whose write is in flight
        self.lengths = {}     # key -> numel
        self.on_disk = set()

    def _paths(self, key):
        return [os.path.join(self.folder, f"{key}_{i}.tensor.swp") for i in range(self.tensors_per_set)]

    def _aligned(self, n):
        return (n + self.align - 1) // self.align * self.align

    def _take_slot(self):
        if not self.free_slots:
            self._finish_writes()
        assert self.free_slots, "all swap buffers are in use: release a state set first"
        return self.free_slots.pop()

    def _finish_writes(self):
        if self.writing:
            self.write_handle.wait()
            self.free_slots.extend(self.writing)
            self.writing = []

    def _finish_read(self):
        if self.reading is not None:
            self.read_handle.wait()
            self.reading = None

    def register(self, key, numel, init_tensors=None):
        """Declare a state set; ``init_tensors`` (or zeros) become its initial on-disk content."""
        self.lengths[key] = numel
        slot = self._take_slot()
        n = self._aligned(numel)
        for i, buf in enumerate(self.slots[slot]):
            view = buf.narrow(0, 0, n)
            view.zero_()
            if init_tensors is not None and init_tensors[i] is not None:
                view.narrow(0, 0, numel).copy_(init_tensors[i].view(-1))
        self._write(key, slot)
        self._finish_writes()

    def prefetch(self, key):
        """Start reading ``key`` (no-op if resident or already in flight)."""
        if key in self.loaded or key not in self.on_disk:
            return
        self._finish_read()
        slot = self._take_slot()
        n = self._aligned(self.lengths[key])
        for buf, path in zip(self.slots[slot], self._paths(key)):
            assert self.read_handle.async_pread(buf.narrow(0, 0, n), path) == 0
        self.loaded[key] = slot
        self.reading = key

    def acquire(self, key):
        """-> list of pinned tensors (length ``numel``) holding the state set; blocks until the read completed."""
        if key not in self.loaded:
            self.prefetch(key)
        if self.reading == key:
            self._finish_read()
        slot = self.loaded[key]
        return [buf.narrow(0, 0, self.lengths[key]) for buf in self.slots[slot]]

    def _write(self, key, slot):
        n = self._aligned(self.lengths[key])
        for buf, path in zip(self.slots[slot], self._paths(key)):
            assert self.write_handle.async_pwrite(buf.narrow(0, 0, n), path) == 0
        self.writing.append(slot)
        self.on_disk.add(key)

    def release(self, key, dirty=True):
        """Give the buffers back; ``dirty`` state sets are written to their files asynchronously first."""
        slot = self.loaded.pop(key)
        if self.reading == key:
            self._finish_read()
        if dirty:
            self._write(key, slot)
        else:
            self.free_slots.append(slot)

    def flush(self):
        self._finish_read()
        self._finish_writes()

    def read_all(self, key):
        """Synchronous copy of a state set (checkpointing)."""
        tensors = [t.clone() for t in self.acquire(key)]
        self.release(key, dirty=False)
        return tensors

    def write_all(self, key, tensors):
        """Overwrite a state set (checkpoint load)."""
        bufs = self.acquire(key) if key in self.on_disk else None
        if bufs is None:
            self.register(key, tensors[0].numel(), tensors)
            return
        for b, t in zip(bufs, tensors):
            b.copy_(t.view(-1))
        self.release(key, dirty=True)
        self.flush()
