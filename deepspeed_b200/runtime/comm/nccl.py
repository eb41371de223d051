"""Reference import path ``deepspeed.runtime.comm.nccl.NcclBackend``."""
from .compressed import CompressedBackend as NcclBackend  # noqa: F401
