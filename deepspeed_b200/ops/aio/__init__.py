"""DeepNVMe async I/O handle (reference ``csrc/aio/py_lib/py_ds_aio.cpp:17-158`` ``aio_handle``)."""
import ctypes
import os

import torch

from ..cpu_lib import lib


class aio_handle:
    """``aio_handle(block_size, queue_depth, single_submit, overlap_events, intra_op_parallelism)`` with
    ``read/write/pread/pwrite/sync_pread/sync_pwrite/async_pread/async_pwrite/wait/new_cpu_locked_tensor``."""

    def __init__(self, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True,
                 intra_op_parallelism=1, use_direct=True):
        self._lib = lib()
        self._h = self._lib.dsb200_aio_create(block_size, queue_depth, max(1, intra_op_parallelism), int(use_direct))
        self._block_size, self._queue_depth = block_size, queue_depth
        self._single_submit, self._overlap_events = single_submit, overlap_events
        self._threads = max(1, intra_op_parallelism)
        self._pinned = {}

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.dsb200_aio_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def get_block_size(self):
        return self._block_size

    def get_queue_depth(self):
        return self._queue_depth

    def get_single_submit(self):
        return self._single_submit

    def get_overlap_events(self):
        return self._overlap_events

    def get_intra_op_parallelism(self):
        return self._threads

    def get_thread_count(self):
        return self._threads

    @staticmethod
    def _check(t):
        assert t.device.type == "cpu" and t.is_contiguous(), "aio operates on contiguous host tensors"

    def pread(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        self._check(buffer)
        rc = self._lib.dsb200_aio_pread(self._h, ctypes.c_void_p(buffer.data_ptr()), buffer.numel() * buffer.element_size(),
                                        os.fsencode(filename), file_offset, int(async_op))
        if rc != 0:
            raise IOError(f"aio pread failed for {filename}")
        return 0 if async_op else 1

    def pwrite(self, buffer, filename, validate=False, async_op=False, file_offset=0):
        self._check(buffer)
        rc = self._lib.dsb200_aio_pwrite(self._h, ctypes.c_void_p(buffer.data_ptr()),
                                         buffer.numel() * buffer.element_size(), os.fsencode(filename), file_offset,
                                         int(async_op))
        if rc != 0:
            raise IOError(f"aio pwrite failed for {filename}")
        return 0 if async_op else 1

    def read(self, buffer, filename, validate=False):
        return self.pread(buffer, filename, validate, False)

    def write(self, buffer, filename, validate=False):
        return self.pwrite(buffer, filename, validate, False)

    def sync_pread(self, buffer, filename, file_offset=0):
        return self.pread(buffer, filename, False, False, file_offset)

    def sync_pwrite(self, buffer, filename, file_offset=0):
        return self.pwrite(buffer, filename, False, False, file_offset)

    def async_pread(self, buffer, filename, file_offset=0):
        return self.pread(buffer, filename, False, True, file_offset)

    def async_pwrite(self, buffer, filename, file_offset=0):
        return self.pwrite(buffer, filename, False, True, file_offset)

    def wait(self):
        n = self._lib.dsb200_aio_wait(self._h)
        if n < 0:
            raise IOError("aio: an I/O request failed")
        return 1

    def new_cpu_locked_tensor(self, num_elem, example_tensor):
        t = torch.empty(num_elem, dtype=example_tensor.dtype, device="cpu",
                        pin_memory=torch.cuda.is_available())
        self._pinned[t.data_ptr()] = t
        return t

    def free_cpu_locked_tensor(self, tensor):
        self._pinned.pop(tensor.data_ptr(), None)
        return True


class AsyncIOBuilder:
    """Shim of the reference's op builder: ``AsyncIOBuilder().load().aio_handle(...)``."""
    NAME = "async_io"

    def is_compatible(self, verbose=False):
        return True

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
