from .base import RolloutBatch, RolloutConfig, RolloutEngine, RolloutRequest, SamplingConfig  # noqa: F401
from .hybrid_engine_rollout import HybridEngineRollout  # noqa: F401

__all__ = ["HybridEngineRollout", "RolloutBatch", "RolloutConfig", "RolloutEngine", "RolloutRequest",
           "SamplingConfig", "build_rollout"]


def build_rollout(rollout_cfg, student_engine=None, tokenizer=None, **kwargs):
    if rollout_cfg.engine == "hybrid_engine":
        if student_engine is None or tokenizer is None:
            raise ValueError("hybrid_engine rollout needs both student_engine and tokenizer")
        return HybridEngineRollout(engine=student_engine, tokenizer=tokenizer, cfg=rollout_cfg)
    raise ValueError(f"Unknown rollout engine {rollout_cfg.engine!r}; choose from 'hybrid_engine'")
