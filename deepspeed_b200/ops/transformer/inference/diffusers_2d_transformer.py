"""Configuration of the diffusers transformer-block replacement (reference
``ops/transformer/inference/diffusers_2d_transformer.py``)."""


class Diffusers2DTransformerConfig:

    def __init__(self, int8_quantization=False):
        self.int8_quantization = int8_quantization
