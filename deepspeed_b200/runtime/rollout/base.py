"""Rollout interface for RL post-training (reference ``runtime/rollout/base.py``)."""
from abc import ABC, abstractmethod
from dataclasses import dataclass

import torch


@dataclass
class RolloutConfig:
    engine: str = "hybrid_engine"
    use_graph_capture: bool = False


@dataclass
class SamplingConfig:
    max_new_tokens: int
    temperature: float = 1.0
    top_p: float = 1.0
    top_k: int = -1
    n_samples_per_prompt: int = 1


@dataclass
class RolloutRequest:
    """Prompts are LEFT padded (real tokens at the right edge) so generation appends after them."""
    prompt_ids: torch.Tensor             # [B, T_p]
    prompt_attention_mask: torch.Tensor  # [B, T_p], 1 on real tokens

    def __post_init__(self):
        if self.prompt_ids.dim() != 2:
            raise ValueError(f"prompt_ids must be 2-D [B, T_p]; got {tuple(self.prompt_ids.shape)}")
        if self.prompt_attention_mask.shape != self.prompt_ids.shape:
            raise ValueError("prompt_attention_mask shape does not match prompt_ids")


@dataclass
class RolloutBatch:
    """Prompt + response, right padded to the longest sequence; B' = B * n_samples_per_prompt."""
    input_ids: torch.Tensor           # [B', T_p + T_r]
    attention_mask: torch.Tensor      # [B', T_p + T_r]
    response_start_idx: torch.Tensor  # [B']

    def __post_init__(self):
        if self.input_ids.dim() != 2:
            raise ValueError(f"input_ids must be 2-D; got {tuple(self.input_ids.shape)}")
        if self.attention_mask.shape != self.input_ids.shape:
            raise ValueError("attention_mask shape does not match input_ids")
        if self.response_start_idx.shape != (self.input_ids.shape[0], ):
            raise ValueError("response_start_idx must be 1-D of length B'")

    @property
    def batch_size(self):
        return int(self.input_ids.shape[0])

    @property
    def seq_len(self):
        return int(self.input_ids.shape[1])


class RolloutEngine(ABC):
    name = "base"

    @abstractmethod
    def generate(self, request: RolloutRequest, sampling: SamplingConfig) -> RolloutBatch:
        ...

    @abstractmethod
    def sync_weights(self, step: int) -> None:
        ...

    def shutdown(self) -> None:
        return None
