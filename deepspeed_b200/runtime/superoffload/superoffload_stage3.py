"""Reference import path; the strategy itself is ``ZeroShardedOptimizer`` with ``offload_optimizer.super_offload``."""
from .superoffload_utils import SuperOffloadCPUOptimizer  # noqa: F401


def _stage3():
    from ..zero.mics import DeepSpeedZeroOptimizer_Stage3
    return DeepSpeedZeroOptimizer_Stage3


class SuperOffloadOptimizer_Stage3:
    """ZeRO-3 with the speculative CPU optimizer under the reference's class name
    (``runtime/superoffload/superoffload_stage3.py``): ``offload_optimizer`` must carry ``super_offload: true``."""

    def __new__(cls, module, optimizer, *, offload_optimizer, **kw):
        assert offload_optimizer is not None and getattr(offload_optimizer, "super_offload", False), \
            "SuperOffloadOptimizer_Stage3 needs zero_optimization.offload_optimizer.super_offload = true"
        return _stage3()(module, optimizer, offload_optimizer=offload_optimizer, **kw)
