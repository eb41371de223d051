from .transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig  # noqa: F401
