"""Compression library: QAT, pruning, layer reduction (reference ``deepspeed/compression``)."""
from .compress import init_compression, redundancy_clean  # noqa: F401
from .scheduler import compression_scheduler  # noqa: F401
from .helper import convert_conv1d_to_linear  # noqa: F401
