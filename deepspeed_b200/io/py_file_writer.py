"""Plain buffered-file writer with the statistics interface of the fast writer (reference ``io/py_file_writer.py``,
``base_file_writer.py``): the baseline the checkpoint benchmarks compare ``FastFileWriter`` against."""
import time

from .base_file_writer import (CLOSE_COUNT_KEY, FLUSH_COUNT_KEY, INVALID_FD, WRITE_BYTES_KEY, WRITE_COUNT_KEY,  # noqa: F401
                               WRITE_SEC_KEY, BaseFileWriter)


class PyFileWriter(BaseFileWriter):

    def __init__(self, file_path):
        super().__init__(file_path)
        self.path = file_path
        self._fp = open(file_path, "wb")

    def write(self, buffer):
        t0 = time.time()
        self._fp.write(buffer)
        self._incr_stats(WRITE_SEC_KEY, time.time() - t0)
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, len(buffer))
        return len(buffer)

    def flush(self):
        self._incr_stats(FLUSH_COUNT_KEY)
        self._fp.flush()

    def close(self):
        self._incr_stats(CLOSE_COUNT_KEY)
        self._fp.close()
