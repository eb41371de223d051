"""Configs of ``deepspeed.linear`` (reference ``linear/config.py``)."""
from dataclasses import dataclass, field
from typing import List

import torch


@dataclass
class LoRAConfig:
    """lora_r: rank; lora_alpha: scaling numerator; base_weight_sharding: shard the frozen base weight over this many
    ranks (= DP world) and all-gather it per forward; offload / offload_ratio: keep (a share of) the frozen weights on
    the host between uses; delay_lora_init: create adapters later via ``init_lora``; target_mods: module names that
    receive adapters."""
    lora_r: int = 64
    lora_alpha: float = 16.
    base_weight_sharding: int = 1
    offload: bool = False
    offload_ratio: float = 0.0
    delay_lora_init: bool = False
    target_mods: List[str] = field(
        default_factory=lambda: ["q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj"])


@dataclass
class QuantizationConfig:
    """q_bits: 8 (e4m3), 6 or 4 (e2m1); mantissa_bits: mantissa width; group_size: elements per scale."""
    q_bits: int = 8
    mantissa_bits: int = 3
    group_size: int = 512
    q_dtype: torch.dtype = torch.uint8
