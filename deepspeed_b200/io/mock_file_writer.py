"""Fake writer for tests of the fast writer path (reference ``deepspeed/io/mock_file_writer.py``)."""


class MockFileWriter:

    def __init__(self, file_path=None):
        self.path = file_path
        self.bytes_written = 0
        self.calls = 0

    def write(self, data):
        self.bytes_written += len(data)
        self.calls += 1
        return len(data)

    def flush(self):
        pass

    def close(self):
        pass
