"""Common face of the checkpoint file writers (reference ``io/base_file_writer.py``, ``io/constants.py``)."""
CLOSE_COUNT_KEY, FILENO_COUNT_KEY, FLUSH_COUNT_KEY = "close", "fileno", "flush"
WRITE_COUNT_KEY, WRITE_BYTES_KEY, WRITE_SEC_KEY, WRITE_SPEED_KEY = "write", "write_bytes", "write_sec", "write_GB/s"
FILE_PATH_KEY = "path"
BASE_STAT_KEYS = [CLOSE_COUNT_KEY, FILENO_COUNT_KEY, FLUSH_COUNT_KEY, WRITE_COUNT_KEY, WRITE_BYTES_KEY, WRITE_SEC_KEY,
                  WRITE_SPEED_KEY]
INVALID_FD = -1


class BaseFileWriter:

    def __init__(self, file_path):
        self._file_path = file_path
        self._stats = {k: 0 for k in BASE_STAT_KEYS}

    def close(self):
        pass

    def fileno(self):
        self._incr_stats(FILENO_COUNT_KEY)
        return INVALID_FD  # torch.save must not bypass write() through the descriptor

    def flush(self):
        pass

    def write(self, buffer):
        pass

    def file_path(self):
        return self._file_path

    def _incr_stats(self, key, incr=1):
        self._stats[key] += incr

    def get_stats(self):
        s = dict(self._stats)
        if s[WRITE_SEC_KEY] > 0:
            s[WRITE_SPEED_KEY] = s[WRITE_BYTES_KEY] / s[WRITE_SEC_KEY] / (1024**3)
        return s

    def _dump_state(self):
        s = self.get_stats()
        s[FILE_PATH_KEY] = self.file_path()
        print(f"stats = {s}")

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
