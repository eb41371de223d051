"""ZenFlow optimizers under the reference's class names (``runtime/zenflow/zenflow_stage_1_and_2.py``).

ZenFlow itself - the important ``topk_ratio`` of every shard is updated on the GPU each step, the rest accumulates and
is updated by the CPU optimizer every ``update_interval`` steps, overlapped with the following steps when
``overlap_step`` is set - is ``zenflow_utils.ZenFlowController`` driven by ``ZeroShardedOptimizer``. The two reference
subclasses differ in exactly that overlap switch."""
import copy


def _base():
    from ..zero.mics import DeepSpeedZeroOptimizer
    return DeepSpeedZeroOptimizer


class ZenFlowZeroOptimizer:
    """Factory: ``ZenFlowZeroOptimizer(module, optimizer, zenflow=cfg, offload_optimizer=..., ...)`` returns the
    stage 1/2 optimizer with the ZenFlow controller attached; ``create(zenflow_config)`` picks the class the
    reference would (sequential or parallel / overlapped)."""
    overlap_step = None

    def __new__(cls, module, optimizer, *, zenflow, offload_optimizer, **kw):
        assert zenflow is not None and offload_optimizer is not None, \
            "ZenFlow needs the zenflow section and zero_optimization.offload_optimizer"
        if cls.overlap_step is not None and getattr(zenflow, "overlap_step", None) != cls.overlap_step:
            zenflow = copy.copy(zenflow)
            zenflow.overlap_step = cls.overlap_step
        return _base()(module, optimizer, zenflow=zenflow, offload_optimizer=offload_optimizer, **kw)

    @classmethod
    def create(cls, zenflow_config):
        return ZenFlowZeroOptimizerParallel if getattr(zenflow_config, "overlap_step", False) \
            else ZenFlowZeroOptimizerSequential


class ZenFlowZeroOptimizerSequential(ZenFlowZeroOptimizer):
    overlap_step = False


class ZenFlowZeroOptimizerParallel(ZenFlowZeroOptimizer):
    overlap_step = True
