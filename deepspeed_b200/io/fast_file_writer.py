"""``FastFileWriter``: file-like object that stages writes in (pinned) double buffers and drains them through the
native async I/O handle (reference ``deepspeed/io/fast_file_writer.py`` used by ``FastCheckpointEngine``).

Data-parallel writing (reference ``num_parallel_writers`` / ``writer_rank``): ranks that hold the same state each
write a share of it. Here the serialised byte stream is cut into fixed-size stripes and stripe ``k`` belongs to writer
``k mod N``; a writer lays its stripes back to back into its own file ``<path>-<rank>.<N>``. No rank needs to know the
stream's total size in advance (the reference partitions each tensor payload and repeats the pickle metadata in every
file), every writer moves ``1/N`` of the bytes, and :class:`StripedReader` presents the ``N`` files as the original
stream again - ``torch.load`` reads through it without the stream ever being assembled in memory."""
import os
import time
import warnings
from dataclasses import dataclass
from typing import Any

import torch

from ..ops.aio import aio_handle
from .base_file_writer import (CLOSE_COUNT_KEY, FLUSH_COUNT_KEY, WRITE_BYTES_KEY, WRITE_COUNT_KEY, WRITE_SEC_KEY,
                               BaseFileWriter)
from .io_buffers import Double_IO_Buffer, Single_IO_Buffer

STRIPE_BYTES = 4 << 20


@dataclass
class FastFileWriterConfig:
    io_buffer_size: int = 1 << 28
    double_buffer: bool = True
    num_threads: int = 4
    block_size: int = 1 << 20
    queue_depth: int = 8
    show_stats: bool = False
    num_parallel_writers: int = 1
    writer_rank: int = 0
    global_rank: int = 0
    stripe_bytes: int = STRIPE_BYTES
    dnvme_handle: Any = None  # reuse a handle / staging buffer across files (the writer factory does)
    pinned_tensor: Any = None


def striped_file_path(file_path, writer_rank, num_writers):
    return f"{file_path}-{writer_rank}.{num_writers}"


class FastFileWriter(BaseFileWriter):

    def __init__(self, file_path, config: FastFileWriterConfig = None):
        self.cfg = config or FastFileWriterConfig()
        c = self.cfg
        assert 0 <= c.writer_rank < c.num_parallel_writers
        super().__init__(file_path)
        self.path = file_path
        self.handle = c.dnvme_handle or aio_handle(c.block_size, c.queue_depth, False, True, c.num_threads)
        buf = c.pinned_tensor
        if buf is None:
            unit = 2 * 4096 if c.double_buffer else 4096
            size = max(unit, (c.io_buffer_size + unit - 1) // unit * unit)
            buf = torch.empty(size, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        self._io = (Double_IO_Buffer if c.double_buffer else Single_IO_Buffer)(buf, self.handle)
        self.file_off = 0  # bytes already handed to the I/O handle
        self.stream_off = 0  # position in the (logical) serialised stream
        self.stats = dict(bytes=0, seconds=0.0)
        open(self.path, "wb").close()
        self._t0 = time.time()

    # ---- stream -> my stripes ------------------------------------------------------------------------------
    def _mine(self, n):
        """Sub-ranges ``(lo, hi)`` of the next ``n`` stream bytes that belong to this writer."""
        N, r, S = self.cfg.num_parallel_writers, self.cfg.writer_rank, self.cfg.stripe_bytes
        if N == 1:
            return [(0, n)]
        out, pos, end = [], self.stream_off, self.stream_off + n
        while pos < end:
            k = pos // S
            stop = min(end, (k + 1) * S)
            if k % N == r:
                out.append((pos - self.stream_off, stop - self.stream_off))
            pos = stop
        return out

    def write(self, data):
        t0 = time.time()
        mv = memoryview(data).cast("B")
        n = len(mv)
        if n:
            with warnings.catch_warnings():  # bytes objects are read-only; they are only read here
                warnings.simplefilter("ignore", UserWarning)
                src = torch.frombuffer(mv, dtype=torch.uint8)
            for lo, hi in self._mine(n):
                pos = lo
                while pos < hi:
                    pos += self._io.fill(src.narrow(0, 0, hi), pos)
                    if self._io.is_full():
                        self._drain()
        self.stream_off += n
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, n)
        self._incr_stats(WRITE_SEC_KEY, time.time() - t0)
        return n

    def _drain(self):
        n = self._io.get_offset()
        if n == 0:
            return
        self._io.drain(n, self.path, self.file_off)
        self.file_off += n
        self.stats["bytes"] += n

    def flush(self):
        self._incr_stats(FLUSH_COUNT_KEY)

    def close(self):
        self._drain()
        self._io.complete_ongoing_drain()
        self._incr_stats(CLOSE_COUNT_KEY)
        self.stats["seconds"] = time.time() - self._t0
        if self.cfg.show_stats:
            gb = self.stats["bytes"] / 2**30
            print(f"FastFileWriter: {gb:.2f} GiB in {self.stats['seconds']:.2f}s = "
                  f"{gb / max(self.stats['seconds'], 1e-9):.2f} GiB/s")


class StripedReader:
    """Read-only file object over the ``N`` files of a striped write (``<path>-<r>.<N>``)."""

    def __init__(self, file_path, num_writers=None, stripe_bytes=STRIPE_BYTES):
        if num_writers is None:
            num_writers = find_striped_writers(file_path)
            if num_writers is None:
                raise FileNotFoundError(f"no striped files '{file_path}-<rank>.<N>'")
        self.N, self.S = num_writers, stripe_bytes
        self._fds = [os.open(striped_file_path(file_path, r, num_writers), os.O_RDONLY) for r in range(num_writers)]
        self._size = sum(os.fstat(fd).st_size for fd in self._fds)
        self._pos = 0

    def readable(self):
        return True

    def seekable(self):
        return True

    def tell(self):
        return self._pos

    def seek(self, offset, whence=0):
        self._pos = offset if whence == 0 else (self._pos + offset if whence == 1 else self._size + offset)
        return self._pos

    def readinto(self, b):
        mv = memoryview(b).cast("B")
        want = min(len(mv), max(0, self._size - self._pos))
        got = 0
        while got < want:
            k, within = divmod(self._pos, self.S)
            r, local_stripe = k % self.N, k // self.N
            n = min(want - got, self.S - within)
            chunk = os.pread(self._fds[r], n, local_stripe * self.S + within)
            if not chunk:
                break
            mv[got:got + len(chunk)] = chunk
            got += len(chunk)
            self._pos += len(chunk)
        return got

    def read(self, n=-1):
        if n is None or n < 0:
            n = self._size - self._pos
        buf = bytearray(min(n, max(0, self._size - self._pos)))
        got = self.readinto(buf)
        return bytes(buf[:got])

    def readline(self, size=-1):  # the unpickler asks for it; lines are short
        out = bytearray()
        while size < 0 or len(out) < size:
            c = self.read(1)
            if not c:
                break
            out += c
            if c == b"\n":
                break
        return bytes(out)

    def close(self):
        for fd in self._fds:
            os.close(fd)
        self._fds = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def find_striped_writers(file_path):
    """``N`` when ``<path>-0.N`` ... ``<path>-(N-1).N`` all exist, else ``None``."""
    d, base = os.path.split(file_path)
    try:
        names = os.listdir(d or ".")
    except FileNotFoundError:
        return None
    for name in names:
        if name.startswith(base + "-0."):
            tail = name[len(base) + 3:]
            if tail.isdigit() and all(os.path.exists(striped_file_path(file_path, r, int(tail)))
                                      for r in range(int(tail))):
                return int(tail)
    return None
