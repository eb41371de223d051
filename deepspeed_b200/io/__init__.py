from .base_file_writer import BaseFileWriter  # noqa: F401
from .io_buffers import Base_IO_Buffer, Double_IO_Buffer, Single_IO_Buffer  # noqa: F401
from .fast_file_writer import (FastFileWriter, FastFileWriterConfig, StripedReader, find_striped_writers,  # noqa: F401
                               striped_file_path)
from .mock_file_writer import MockFileWriter  # noqa: F401
from .py_file_writer import PyFileWriter  # noqa: F401
from .utils import bytes_to_tensor, obj_serialization_details, required_minimum_torch_version, tensor_to_bytes  # noqa: F401,E402
