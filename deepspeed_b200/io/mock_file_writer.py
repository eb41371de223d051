"""Fake writer for tests of the fast writer path (reference ``deepspeed/io/mock_file_writer.py``)."""
from .base_file_writer import WRITE_BYTES_KEY, WRITE_COUNT_KEY, BaseFileWriter


class MockFileWriter(BaseFileWriter):

    def __init__(self, file_path=None):
        super().__init__(file_path)
        self.path = file_path
        self.bytes_written = 0
        self.calls = 0

    def write(self, data):
        self.bytes_written += len(data)
        self.calls += 1
        self._incr_stats(WRITE_COUNT_KEY)
        self._incr_stats(WRITE_BYTES_KEY, len(data))
        return len(data)
