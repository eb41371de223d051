"""Parallelism strategies: AutoTP (tp), AutoEP (auto_ep), sequence parallel lives in ``deepspeed_b200.sequence``.
Also the import surface of the reference's ``deepspeed.module_inject``."""
from .tp import (AutoTP, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, VocabParallelEmbedding, apply_autotp,
                 shard_tensor)
from .layers import (AsyncColumnParallel, ColumnParallel, Conv_LinearALlreduce, EmbeddingLayer,  # noqa: F401
                     FusedModuleWrapper, GateUpPack_LinearLayer, GatherFromTensorParallelRegion,
                     GatherReplacedLayerParams, Normalize, OPTEmbedding, RMSNormalize, RowParallel,
                     SubParamColumnParallel, SubParamLinearAllreduce, SubParamLinearLayer, TensorParallel_Layer,
                     TensorParallelConv2d, TensorParallelIcShardConv2d, TensorParallelOcShardConv2d,
                     Yuan_LinearAllreduce, Yuan_LinearLayer, conv_LinearLayer, fused_LinearLayer,
                     is_autotp_training_mode, set_autotp_mode)
from .autotp_config import AutoTPConfig, AutoTPPresets, PartitionType, TPLayerSpec, merge_autotp_configs  # noqa: F401
from .replace_module import (DSPolicy, GroupQuantizer, HFBertLayerPolicy, ReplaceWithTensorSlicing,  # noqa: F401
                             generic_injection, quantize_transformer_layer, replace_transformer_layer,
                             revert_transformer_layer)
from .auto_ep import AutoEP, AutoEPMoELayer, TokenChoiceTopKRouter, apply_autoep  # noqa: F401,E402
from .auto_ep_presets import AutoEPConfig, MoELayerSpec, MoEModelPreset, parse_autoep_config  # noqa: F401,E402
from .auto_ep_folding import ParallelFoldingSpec, build_folding_spec  # noqa: F401,E402
