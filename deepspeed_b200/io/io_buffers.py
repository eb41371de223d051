"""Staging buffers between ``write()`` calls and the native async writer (reference ``io/base_io_buffer.py``,
``single_io_buffer.py``, ``double_io_buffer.py``).

A buffer takes bytes until it is full, then ``drain`` hands the filled part to the async I/O handle as one positional
write. The double buffer keeps filling its second half while the first is still in flight: serialisation (host) and the
file write (I/O threads) overlap."""
import torch


class Base_IO_Buffer:

    def __init__(self, pinned_tensor, dnvme_handle):
        assert pinned_tensor.numel() % 4096 == 0, "the staging buffer must be a multiple of the 4 KiB I/O alignment"
        self._dnvme_handle = dnvme_handle
        self._pinned_tensor = pinned_tensor

    def fill(self, src_tensor, src_offset):
        raise NotImplementedError

    def drain(self, num_bytes, fd_or_path, file_offset):
        raise NotImplementedError

    def is_empty(self):
        return self.get_offset() == 0

    def is_full(self):
        raise NotImplementedError

    def get_buffer(self):
        raise NotImplementedError

    def get_offset(self):
        raise NotImplementedError

    def get_aligned_num_bytes(self):
        return self.get_offset() // 4096 * 4096

    def get_unaligned_num_bytes(self):
        return self.get_offset() % 4096

    def reset(self):
        raise NotImplementedError

    def complete_ongoing_drain(self):
        raise NotImplementedError

    def _copy_in(self, dst, dst_offset, src_tensor, src_offset):
        n = min(src_tensor.numel() - src_offset, dst.numel() - dst_offset)
        dst.narrow(0, dst_offset, n).copy_(src_tensor.narrow(0, src_offset, n))
        return n


class Single_IO_Buffer(Base_IO_Buffer):

    def __init__(self, pinned_tensor, dnvme_handle):
        super().__init__(pinned_tensor, dnvme_handle)
        self._offset = 0

    def fill(self, src_tensor, src_offset):
        n = self._copy_in(self._pinned_tensor, self._offset, src_tensor, src_offset)
        self._offset += n
        return n

    def drain(self, num_bytes, fd_or_path, file_offset):
        self._dnvme_handle.sync_pwrite(self._pinned_tensor.narrow(0, 0, num_bytes), fd_or_path, file_offset)
        self._offset = 0

    def is_full(self):
        return self._offset == self._pinned_tensor.numel()

    def get_buffer(self):
        return self._pinned_tensor

    def get_offset(self):
        return self._offset

    def reset(self):
        self._offset = 0

    def complete_ongoing_drain(self):
        pass


class Double_IO_Buffer(Base_IO_Buffer):

    def __init__(self, pinned_tensor, dnvme_handle):
        super().__init__(pinned_tensor, dnvme_handle)
        assert pinned_tensor.numel() % (2 * 4096) == 0
        half = pinned_tensor.numel() // 2
        self._halves = [pinned_tensor.narrow(0, 0, half), pinned_tensor.narrow(0, half, half)]
        self._offsets = [0, 0]
        self._index = 0
        self._inflight = False

    def fill(self, src_tensor, src_offset):
        n = self._copy_in(self._halves[self._index], self._offsets[self._index], src_tensor, src_offset)
        self._offsets[self._index] += n
        return n

    def drain(self, num_bytes, fd_or_path, file_offset):
        self.complete_ongoing_drain()
        self._dnvme_handle.async_pwrite(self._halves[self._index].narrow(0, 0, num_bytes), fd_or_path, file_offset)
        self._inflight = True
        self._index = 1 - self._index  # the other half (drained one call ago, waited for above) takes the next bytes
        self._offsets[self._index] = 0

    def is_full(self):
        return self._offsets[self._index] == self._halves[self._index].numel()

    def get_buffer(self):
        return self._halves[self._index]

    def get_offset(self):
        return self._offsets[self._index]

    def reset(self):
        self._offsets[self._index] = 0

    def complete_ongoing_drain(self):
        if self._inflight:
            self._dnvme_handle.wait()
            self._inflight = False
