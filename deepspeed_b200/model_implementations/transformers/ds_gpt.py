from .ds_transformer import DeepSpeedTransformerInference


class DeepSpeedGPTInference(DeepSpeedTransformerInference):
    """GPT flavour of the v1 inference layer: the differences between the families are configuration
    (``DeepSpeedInferenceConfig``: norm placement, rotary, alibi / BLOOM layout, activation), not code."""

    def __init__(self, config, mp_group=None, quantize_scales=None, quantize_groups=1, merge_count=1,
                 mlp_extra_grouping=False):
        super().__init__(config, mp_group, quantize_scales, quantize_groups, merge_count, mlp_extra_grouping)
