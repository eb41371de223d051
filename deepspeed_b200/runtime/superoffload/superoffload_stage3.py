"""Reference import path; the strategy itself is ``ZeroShardedOptimizer`` with ``offload_optimizer.super_offload``."""
from .superoffload_utils import SuperOffloadCPUOptimizer  # noqa: F401
