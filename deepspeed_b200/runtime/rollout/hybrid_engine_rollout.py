"""Rollouts through the training engine itself (reference ``runtime/rollout/hybrid_engine_rollout.py:29``): weights are
shared, so ``sync_weights`` is a no-op. Decoding keeps static shapes (pre-allocated [B', T_p + T_new] buffers, one
new column per step) so the loop is CUDA-graph friendly; sampling is temperature / top-k / top-p."""
from dataclasses import dataclass

import torch

from ..hybrid_engine import sample_top_p
from .base import RolloutBatch, RolloutConfig, RolloutEngine, RolloutRequest, SamplingConfig


@dataclass
class HybridEngineRolloutConfig(RolloutConfig):
    engine: str = "hybrid_engine"


class HybridEngineRollout(RolloutEngine):
    name = "hybrid_engine"

    def __init__(self, engine, tokenizer, cfg=None):
        self.engine, self.tokenizer = engine, tokenizer
        self.cfg = cfg or HybridEngineRolloutConfig()

    def _pad_id(self):
        pid = getattr(self.tokenizer, "pad_token_id", None)
        return pid if pid is not None else (getattr(self.tokenizer, "eos_token_id", 0) or 0)

    @torch.no_grad()
    def generate(self, request: RolloutRequest, sampling: SamplingConfig) -> RolloutBatch:
        n = max(1, sampling.n_samples_per_prompt)
        ids = request.prompt_ids.repeat_interleave(n, dim=0)
        mask = request.prompt_attention_mask.repeat_interleave(n, dim=0)
        B, Tp = ids.shape
        T = Tp + sampling.max_new_tokens
        pad, eos = self._pad_id(), getattr(self.tokenizer, "eos_token_id", None)
        out_ids = torch.full((B, T), pad, dtype=ids.dtype, device=ids.device)
        out_mask = torch.zeros((B, T), dtype=mask.dtype, device=ids.device)
        out_ids[:, :Tp], out_mask[:, :Tp] = ids, mask
        finished = torch.zeros(B, dtype=torch.bool, device=ids.device)
        module = self.engine.module
        was_training = module.training
        module.eval()
        hybrid = hasattr(self.engine, "_enter_generation")
        if hybrid:
            self.engine._enter_generation()
        try:
            for t in range(Tp, T):
                logits = self._forward(module, out_ids[:, :t], out_mask[:, :t])[:, -1, :].float()
                if sampling.top_k and sampling.top_k > 0:
                    kth = logits.topk(sampling.top_k, dim=-1).values[:, -1:]
                    logits = logits.masked_fill(logits < kth, float("-inf"))
                if sampling.temperature <= 0:
                    nxt = logits.argmax(-1)
                else:
                    nxt = sample_top_p(logits, sampling.temperature, sampling.top_p)
                nxt = torch.where(finished, torch.full_like(nxt, pad), nxt)
                out_ids[:, t] = nxt
                out_mask[:, t] = (~finished).to(out_mask.dtype)
                if eos is not None:
                    finished |= nxt == eos
                    if bool(finished.all()):
                        break
        finally:
            if hybrid:
                self.engine._exit_generation()
            if was_training:
                module.train()
        start = torch.full((B, ), Tp, dtype=torch.long, device=ids.device)
        return RolloutBatch(input_ids=out_ids, attention_mask=out_mask, response_start_idx=start)

    @staticmethod
    def _forward(module, ids, mask):
        try:
            out = module(input_ids=ids, attention_mask=mask)
        except TypeError:
            out = module(ids)
        return out.logits if hasattr(out, "logits") else (out[0] if isinstance(out, (tuple, list)) else out)

    _sample_top_p = staticmethod(sample_top_p)

    def sync_weights(self, step: int) -> None:
        return None
