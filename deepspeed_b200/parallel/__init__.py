"""Parallelism strategies: AutoTP (tp), AutoEP (auto_ep), sequence parallel lives in ``deepspeed_b200.sequence``.
Also the import surface of the reference's ``deepspeed.module_inject``."""
from .tp import (AutoTP, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, VocabParallelEmbedding, apply_autotp,
                 shard_tensor)
from .layers import (EmbeddingLayer, Normalize, SubParamLinearAllreduce, SubParamLinearLayer, set_autotp_mode,  # noqa: F401
                     is_autotp_training_mode)
from .autotp_config import AutoTPConfig, AutoTPPresets, PartitionType, TPLayerSpec, merge_autotp_configs  # noqa: F401
from .replace_module import (DSPolicy, GroupQuantizer, HFBertLayerPolicy, ReplaceWithTensorSlicing,  # noqa: F401
                             generic_injection, quantize_transformer_layer, replace_transformer_layer,
                             revert_transformer_layer)
from .auto_ep import AutoEP, AutoEPMoELayer, TokenChoiceTopKRouter, apply_autoep  # noqa: F401,E402
from .auto_ep_presets import AutoEPConfig, MoELayerSpec, MoEModelPreset, parse_autoep_config  # noqa: F401,E402
from .auto_ep_folding import ParallelFoldingSpec, build_folding_spec  # noqa: F401,E402
