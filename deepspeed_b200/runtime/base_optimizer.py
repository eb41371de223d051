"""Marker base classes of the optimizer wrappers (reference ``runtime/base_optimizer.py``): user code tests
``isinstance(engine.optimizer, DeepSpeedOptimizer)`` / ``ZeROOptimizer``."""


class DeepSpeedOptimizer(object):
    pass


class ZeROOptimizer(DeepSpeedOptimizer):
    """Base of the ZeRO optimizers; here there is one implementation for stages 0-3 (``ZeroShardedOptimizer``)."""
    pass
