from ...ops.transformer.inference.ds_transformer import (BloomSelfAttention, DeepSpeedMLP, DeepSpeedSelfAttention,  # noqa: F401
                                                         DeepSpeedTransformerInference)
