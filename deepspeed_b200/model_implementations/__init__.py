"""Reference import path ``deepspeed.model_implementations``: the v1 inference layer classes and the CUDA-graph wrappers
for diffusers sub-models."""
from .transformers.ds_transformer import DeepSpeedTransformerInference  # noqa: F401
from .transformers.clip_encoder import DSClipEncoder  # noqa: F401
