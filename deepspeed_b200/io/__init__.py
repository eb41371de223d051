from .fast_file_writer import FastFileWriter, FastFileWriterConfig
from .mock_file_writer import MockFileWriter
