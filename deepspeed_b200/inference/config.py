"""Inference config (reference ``inference/config.py:118`` ``DeepSpeedInferenceConfig``)."""
from typing import Any, Dict, Optional, Union

import torch
from pydantic import Field, field_validator

from ..config.base import DeepSpeedConfigModel


from enum import Enum  # noqa: E402


class DtypeEnum(Enum):
    """Accepted spellings of the inference dtypes (reference ``inference/config.py:15``)."""
    fp16 = (torch.float16, "torch.float16", "fp16", "float16", "half")
    fp32 = (torch.float32, "torch.float32", "fp32", "float32", "float")
    bf16 = (torch.bfloat16, "torch.bfloat16", "bf16", "bfloat16", "bfloat")
    int8 = (torch.int8, "torch.int8", "int8")

    @classmethod
    def from_str(cls, value):
        for dtype in cls:
            if value in dtype.value:
                return dtype
        raise ValueError(f"'{value}' is not a valid DtypeEnum")


class MoETypeEnum(str, Enum):
    residual = "residual"
    standard = "standard"


class QuantTypeEnum(str, Enum):
    asym = "asymmetric"
    sym = "symmetric"


class BaseQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True
    num_bits: int = 8
    q_type: QuantTypeEnum = QuantTypeEnum.sym
    q_groups: int = 1


class WeightQuantConfig(BaseQuantConfig):
    enabled: bool = True
    quantized_initialization: Dict = Field(default_factory=dict)
    post_init_quant: Dict = Field(default_factory=dict)


class ActivationQuantConfig(BaseQuantConfig):
    enabled: bool = True


class QKVQuantConfig(DeepSpeedConfigModel):
    enabled: bool = True


class DeepSpeedTPConfig(DeepSpeedConfigModel):
    enabled: bool = True
    tp_size: int = 1
    tp_grain_size: int = 64
    mpu: Any = None
    tp_group: Any = None


class DeepSpeedMoEConfig(DeepSpeedConfigModel):
    enabled: bool = True
    ep_size: int = 1
    moe_experts: list = Field([1], alias="num_experts")
    type: str = "standard"
    ep_mp_group: Any = None
    ep_group: Any = None


class QuantizationConfig(DeepSpeedConfigModel):
    enabled: bool = True
    activation: Dict[str, Any] = Field(default_factory=dict)
    weight: Dict[str, Any] = Field(default_factory=dict)
    qkv: Dict[str, Any] = Field(default_factory=dict)


class InferenceCheckpointConfig(DeepSpeedConfigModel):
    checkpoint_dir: Optional[str] = None
    save_mp_checkpoint_path: Optional[str] = None
    base_dir: Optional[str] = None


class DeepSpeedInferenceConfig(DeepSpeedConfigModel):
    replace_with_kernel_inject: bool = Field(False, alias="kernel_inject")
    dtype: Any = torch.float16
    tensor_parallel: DeepSpeedTPConfig = Field(default_factory=DeepSpeedTPConfig, alias="tp")
    enable_cuda_graph: bool = False
    use_triton: bool = False
    triton_autotune: bool = False
    zero: Dict[str, Any] = Field(default_factory=dict)
    triangular_masking: bool = Field(True, alias="tm")
    moe: Union[bool, DeepSpeedMoEConfig] = Field(default_factory=DeepSpeedMoEConfig)
    keep_module_on_host: bool = False
    quant: QuantizationConfig = Field(default_factory=QuantizationConfig)
    checkpoint: Optional[Union[str, Dict]] = None
    base_dir: str = ""
    set_empty_params: bool = False
    save_mp_checkpoint_path: Optional[str] = None
    checkpoint_config: InferenceCheckpointConfig = Field(default_factory=InferenceCheckpointConfig, alias="ckpt_config")
    return_tuple: bool = True
    training_mp_size: int = 1
    replace_method: str = "auto"
    injection_policy: Optional[Dict] = Field(None, alias="injection_dict")
    injection_policy_tuple: Optional[tuple] = None
    config: Optional[Dict] = Field(None, alias="args")
    max_out_tokens: int = Field(1024, alias="max_tokens")
    min_out_tokens: int = Field(1, alias="min_tokens")
    transposed_mode: bool = False
    mp_size: int = Field(1, json_schema_extra={"deprecated": True, "new_param": "tensor_parallel.tp_size"})
    mpu: Any = None
    ep_size: int = 1
    ep_group: Any = None
    ep_mp_group: Any = None
    moe_experts: list = [1]
    moe_type: str = "standard"

    @field_validator("dtype", mode="before")
    @classmethod
    def _dtype(cls, v):
        if isinstance(v, str):
            table = {"fp16": torch.float16, "half": torch.float16, "float16": torch.float16, "bf16": torch.bfloat16,
                     "bfloat16": torch.bfloat16, "fp32": torch.float32, "float32": torch.float32, "float": torch.float32,
                     "int8": torch.int8, "torch.float16": torch.float16, "torch.bfloat16": torch.bfloat16,
                     "torch.float32": torch.float32}
            if v not in table:
                raise ValueError(f"unknown dtype {v}")
            return table[v]
        return v
