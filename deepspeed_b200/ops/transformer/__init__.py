from .transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig  # noqa: F401
from .inference import (DeepSpeedInferenceConfig, DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig,  # noqa: F401
                        DeepSpeedTransformerInference)
