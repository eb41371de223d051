"""Inference: v1 engine (AutoTP + CUDA graphs) and v2 ragged-batching engine with a blocked KV cache."""
from .config import DeepSpeedInferenceConfig
from .engine import InferenceEngine
from .v2 import (InferenceEngineV2, RaggedInferenceEngineConfig, DeepSpeedTPConfig, build_hf_engine,  # noqa: F401,E402
                 build_engine_from_ds_checkpoint)
