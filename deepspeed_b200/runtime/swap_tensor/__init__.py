"""ZeRO-Infinity tensor swapping between pinned host memory and NVMe files through the native async I/O library
(``csrc/cpu/aio.cpp``, reference ``runtime/swap_tensor/*`` over ``csrc/aio``)."""
from .aio_config import get_aio_config  # noqa: F401
from .async_swapper import AsyncTensorSwapper  # noqa: F401
from .optimizer_utils import OptimizerSwapper  # noqa: F401
from .partitioned_optimizer_swapper import PartitionedOptimizerSwapper  # noqa: F401
from .pipelined_optimizer_swapper import PipelinedOptimizerSwapper  # noqa: F401
from .partitioned_param_swapper import AsyncPartitionedParameterSwapper, PartitionedParamStatus  # noqa: F401

MIN_AIO_BYTES = 1024**2
MIN_SWAPPABLE_BYTES = MIN_AIO_BYTES  # tensors below this stay in memory (reference swap_tensor/utils.py:16)
