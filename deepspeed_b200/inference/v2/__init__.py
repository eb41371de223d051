from .engine_v2 import InferenceEngineV2, RaggedInferenceEngineConfig, greedy_generate
from .ragged import BlockedAllocator, BlockedKVCache, DSSequenceDescriptor, DSStateManager, RaggedBatchWrapper
