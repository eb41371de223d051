"""``FastFileWriter``: file-like object that stages writes in (pinned) double buffers and drains them through the
native async I/O handle (reference ``deepspeed/io/fast_file_writer.py`` used by ``FastCheckpointEngine``)."""
import os
import time
from dataclasses import dataclass

import torch

from ..ops.aio import aio_handle


@dataclass
class FastFileWriterConfig:
    io_buffer_size: int = 1 << 28
    double_buffer: bool = True
    num_threads: int = 4
    block_size: int = 1 << 20
    queue_depth: int = 8
    show_stats: bool = False


class FastFileWriter:

    def __init__(self, file_path, config: FastFileWriterConfig = None):
        self.cfg = config or FastFileWriterConfig()
        self.path = file_path
        n = 2 if self.cfg.double_buffer else 1
        size = (self.cfg.io_buffer_size + 4095) // 4096 * 4096
        self.buffers = [torch.empty(size, dtype=torch.uint8, pin_memory=torch.cuda.is_available()) for _ in range(n)]
        self.handle = aio_handle(self.cfg.block_size, self.cfg.queue_depth, False, True, self.cfg.num_threads)
        self.cur, self.fill, self.file_off = 0, 0, 0
        self.pending = False
        self.stats = dict(bytes=0, seconds=0.0)
        open(self.path, "wb").close()
        self._t0 = time.time()

    def write(self, data):
        mv = memoryview(data).cast("B")
        n, pos = len(mv), 0
        while pos < n:
            buf = self.buffers[self.cur]
            k = min(n - pos, buf.numel() - self.fill)
            buf[self.fill:self.fill + k] = torch.frombuffer(mv[pos:pos + k], dtype=torch.uint8)
            self.fill += k
            pos += k
            if self.fill == buf.numel():
                self._drain()
        return n

    def _drain(self, final=False):
        if self.fill == 0:
            return
        if self.pending:
            self.handle.wait()
            self.pending = False
        buf = self.buffers[self.cur][:self.fill]
        self.handle.async_pwrite(buf, self.path, self.file_off)
        self.pending = True
        self.file_off += self.fill
        self.stats["bytes"] += self.fill
        self.fill = 0
        if len(self.buffers) > 1:
            self.cur = 1 - self.cur
        else:
            self.handle.wait()
            self.pending = False

    def flush(self):
        pass

    def close(self):
        self._drain(final=True)
        if self.pending:
            self.handle.wait()
            self.pending = False
        self.stats["seconds"] = time.time() - self._t0
        if self.cfg.show_stats:
            gb = self.stats["bytes"] / 2**30
            print(f"FastFileWriter: {gb:.2f} GiB in {self.stats['seconds']:.2f}s = {gb / max(self.stats['seconds'], 1e-9):.2f} GiB/s")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
