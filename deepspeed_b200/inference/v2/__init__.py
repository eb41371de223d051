from .engine_v2 import InferenceEngineV2, RaggedInferenceEngineConfig, greedy_generate
from .ragged import BlockedAllocator, BlockedKVCache, DSSequenceDescriptor, DSStateManager, RaggedBatchWrapper
from .engine_factory import DeepSpeedTPConfig, build_engine_from_ds_checkpoint, build_hf_engine  # noqa: F401,E402
