"""Reference import path ``deepspeed.pipe``."""
from ..runtime.pipe import LayerSpec, PipelineModule, TiedLayerSpec  # noqa: F401
