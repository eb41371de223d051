"""Checkpoint writer selection and data-parallel write sharing (reference ``runtime/model_checkpointing``)."""
from .config import get_checkpoint_config, get_checkpoint_writer_config  # noqa: F401
from .constants import *  # noqa: F401,F403
from .data_parallel_writer_factory import DataParallelWriterConfig, DataParallelWriterFactory  # noqa: F401
from .utils import (ExpertParallelInfo, MPUInfo, UniversalParallelInfo, create_data_parallel_writer_config,  # noqa: F401
                    create_universal_parallel_info)
from .writer_factory import CheckpointWriterFactory  # noqa: F401
