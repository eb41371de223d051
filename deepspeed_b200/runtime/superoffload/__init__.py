from .superoffload_utils import SuperOffloadCPUOptimizer  # noqa: F401
