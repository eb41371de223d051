from .fast_file_writer import FastFileWriter, FastFileWriterConfig
from .mock_file_writer import MockFileWriter
from .py_file_writer import PyFileWriter  # noqa: F401,E402
