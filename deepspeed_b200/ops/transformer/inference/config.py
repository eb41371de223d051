"""Configuration of the v1 inference transformer layer (reference ``ops/transformer/inference/config.py``)."""
import json

import torch

from ....utils.types import ActivationFuncType, NormType


class TransformerConfig:

    def __init__(self, hidden_size, intermediate_size, heads, num_hidden_layers):
        self.layer_id = -1
        self.hidden_size = hidden_size
        self.intermediate_size = intermediate_size
        self.heads = heads
        self.num_hidden_layers = num_hidden_layers


class DeepSpeedInferenceConfig(TransformerConfig):
    """Every switch of the reference layer: pre/post layer norm, norm type, rotary (dim, half / interleaved, theta),
    causal (``triangular_masking``) and local-window attention, activation, parallel attention + MLP
    (``mlp_after_attn=False``), GQA (``num_kv``), BLOOM alibi layout, mask inversion, return conventions."""

    def __init__(self, hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12,
                 local_rank=-1, mp_size=1, dtype=torch.float16, pre_layer_norm=True, norm_type=NormType.LayerNorm,
                 stochastic_mode=False, scale_attention=True, triangular_masking=True, local_attention=False,
                 window_size=256, rotary_dim=-1, rotate_half=False, rotate_every_two=True, return_tuple=True,
                 mlp_after_attn=True, mlp_act_func_type=ActivationFuncType.GELU, training_mp_size=1,
                 bigscience_bloom=False, max_out_tokens=1024, min_out_tokens=1, enable_qkv_quantization=False,
                 use_mup=False, scale_attn_by_inverse_layer_idx=False, return_single_tuple=False,
                 set_empty_params=False, transposed_mode=False, use_triton=False, triton_autotune=False, num_kv=-1,
                 rope_theta=10000, invert_mask=True):
        super().__init__(hidden_size, intermediate_size if intermediate_size > 0 else 4 * hidden_size, heads,
                         num_hidden_layers)
        self.dtype, self.pre_layer_norm, self.norm_type, self.local_rank = dtype, pre_layer_norm, norm_type, local_rank
        self.stochastic_mode, self.epsilon, self.mp_size = stochastic_mode, layer_norm_eps, mp_size
        self.scale_attention, self.triangular_masking = scale_attention, triangular_masking
        self.local_attention, self.window_size = local_attention, window_size
        self.rotary_dim, self.rotate_half, self.rotate_every_two = rotary_dim, rotate_half, rotate_every_two
        self.return_tuple, self.mlp_after_attn, self.mlp_act_func_type = return_tuple, mlp_after_attn, mlp_act_func_type
        self.training_mp_size, self.bigscience_bloom = training_mp_size, bigscience_bloom
        self.max_out_tokens, self.min_out_tokens = max_out_tokens, min_out_tokens
        self.enable_qkv_quantization, self.use_mup = enable_qkv_quantization, use_mup
        self.scale_attn_by_inverse_layer_idx, self.return_single_tuple = scale_attn_by_inverse_layer_idx, return_single_tuple
        self.set_empty_params, self.transposed_mode = set_empty_params, transposed_mode
        self.use_triton, self.triton_autotune = False, False  # no Triton in this framework
        self.num_kv, self.rope_theta, self.invert_mask = num_kv, rope_theta, invert_mask

    @classmethod
    def from_dict(cls, json_object):
        config = cls()
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-8") as reader:
            return cls.from_dict(json.loads(reader.read()))
