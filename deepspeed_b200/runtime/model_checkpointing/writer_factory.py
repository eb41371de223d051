"""Builds the file writer a checkpoint engine serialises into (reference
``runtime/model_checkpointing/writer_factory.py``): mock / plain python / fast (pinned staging buffers + native async
I/O). One staging buffer and one I/O handle are kept for all files of the factory; with ``optimize_dp_state`` the fast
writer stripes the stream over the elected data-parallel writers (``io/fast_file_writer.py``)."""
import torch

from ...io import FastFileWriter, FastFileWriterConfig, MockFileWriter, PyFileWriter, striped_file_path
from .constants import (CHECKPOINT_IO_BUFFER_DOUBLE, CHECKPOINT_IO_BUFFER_SIZE, CHECKPOINT_IO_MULTIPLIER,
                        CHECKPOINT_IO_STATISTICS, CHECKPOINT_WRITER_TYPE, CheckpointWriterType)


class CheckpointWriterFactory:

    def __init__(self, writer_config, aio_config, dp_writer_config):
        self._type = writer_config[CHECKPOINT_WRITER_TYPE]
        self._io_buffer_size = writer_config[CHECKPOINT_IO_BUFFER_SIZE]
        self._io_buffer_double = writer_config[CHECKPOINT_IO_BUFFER_DOUBLE]
        self._io_multiplier = writer_config.get(CHECKPOINT_IO_MULTIPLIER, 1)
        self._show_statistics = bool(writer_config.get(CHECKPOINT_IO_STATISTICS, False))
        self._data_parallel_writer = dp_writer_config
        self._aio = aio_config
        self._io_buffer = None
        self._dnvme_handle = None
        self._writer = None
        if self._type == CheckpointWriterType.FAST:
            self._setup_for_aio()

    def _aio_get(self, key, default):
        a = self._aio
        if a is None:
            return default
        return a.get(key, default) if isinstance(a, dict) else getattr(a, key, default)

    def _setup_for_aio(self):
        from ...ops.aio import aio_handle
        unit = 2 * 4096 if self._io_buffer_double else 4096
        size = max(unit, (int(self._io_buffer_size) + unit - 1) // unit * unit)
        self._io_buffer = torch.zeros(size, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        self._dnvme_handle = aio_handle(self._aio_get("block_size", 1 << 20), self._aio_get("queue_depth", 8),
                                        self._aio_get("single_submit", False), self._aio_get("overlap_events", True),
                                        self._aio_get("intra_op_parallelism", 4))

    def create_writer(self, file_path, optimize_dp_state):
        assert self._writer is None, (f"Cannot create checkpoint writer for {file_path}: the writer is still used for "
                                      f"{self._writer.file_path()}; call release_writer() first")
        if self._type == CheckpointWriterType.MOCK:
            self._writer = MockFileWriter(file_path)
        elif self._type == CheckpointWriterType.PYTHON:
            self._writer = PyFileWriter(file_path)
        else:
            dp = self._data_parallel_writer
            n, r = (dp.world_size, dp.rank) if optimize_dp_state and dp is not None else (1, 0)
            if n > 1:
                file_path = striped_file_path(file_path, r, n)
            cfg = FastFileWriterConfig(io_buffer_size=self._io_buffer.numel(), double_buffer=self._io_buffer_double,
                                       show_stats=False, num_parallel_writers=n, writer_rank=r,
                                       global_rank=dp.global_rank if dp is not None else 0,
                                       dnvme_handle=self._dnvme_handle, pinned_tensor=self._io_buffer)
            self._writer = FastFileWriter(file_path, cfg)
        return self._writer

    def release_writer(self):
        self._writer.close()
        if self._show_statistics:
            self._writer._dump_state()
        self._writer = None
