from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ds_transformer import (BloomSelfAttention, DeepSpeedMLP, DeepSpeedSelfAttention,  # noqa: F401
                             DeepSpeedTransformerInference)
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig  # noqa: F401
