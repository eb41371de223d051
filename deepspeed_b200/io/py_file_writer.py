"""Plain buffered-file writer with the statistics interface of the fast writer (reference ``io/py_file_writer.py``,
``base_file_writer.py``): the baseline the checkpoint benchmarks compare ``FastFileWriter`` against."""
import time

INVALID_FD = -1


class PyFileWriter:

    def __init__(self, file_path):
        self.path = file_path
        self._fp = open(file_path, "wb")
        self._stats = {"close": 0, "fileno": 0, "flush": 0, "write": 0, "write_bytes": 0, "write_sec": 0.0}

    def write(self, buffer):
        t0 = time.time()
        self._fp.write(buffer)
        self._stats["write_sec"] += time.time() - t0
        self._stats["write"] += 1
        self._stats["write_bytes"] += len(buffer)
        return len(buffer)

    def flush(self):
        self._stats["flush"] += 1
        self._fp.flush()

    def fileno(self):
        self._stats["fileno"] += 1
        return INVALID_FD  # torch.save must not bypass write() through the descriptor

    def close(self):
        self._stats["close"] += 1
        self._fp.close()

    def get_stats(self):
        return dict(self._stats)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
