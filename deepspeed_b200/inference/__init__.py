"""Inference: v1 engine (AutoTP + CUDA graphs) and v2 ragged-batching engine with a blocked KV cache."""
from .config import DeepSpeedInferenceConfig
from .engine import InferenceEngine
