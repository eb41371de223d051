"""``FP16_Optimizer`` (reference ``runtime/fp16/fused_optimizer.py:33``): fp16 parameters, flat fp32 master copy, static
or dynamic loss scaling with overflow skip - here ``ZeroShardedOptimizer(stage=0, dtype=fp16)``."""
from types import SimpleNamespace

import torch

from ..zero.sharded_optimizer import ZeroShardedOptimizer


def _fp16_cfg(static_loss_scale, dynamic_loss_scale, dynamic_loss_args, initial_dynamic_scale):
    args = dict(init_scale=initial_dynamic_scale, scale_window=1000, min_scale=1, delayed_shift=2,
                consecutive_hysteresis=False)
    args.update(dynamic_loss_args or {})
    return SimpleNamespace(loss_scale=0 if dynamic_loss_scale else static_loss_scale,
                           dynamic_loss_scale_args=lambda: args)


class FP16_Optimizer(ZeroShardedOptimizer):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 initial_dynamic_scale=2**32, dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0,
                 fused_adam_legacy=False, has_moe_layers=False, timers=None, module=None):
        from ... import comm as dist
        module = module if module is not None else getattr(deepspeed, "module", None)
        assert module is not None, "FP16_Optimizer needs the module (pass deepspeed=engine or module=...)"
        world = dist.get_world_size() if dist.is_initialized() else 1
        super().__init__(module, init_optimizer, stage=0, dp_group=None, dp_world=world,
                         dp_rank=dist.get_rank() if dist.is_initialized() else 0,
                         mp_group=(mpu.get_model_parallel_group() if mpu is not None else None), clip_grad=clip_grad,
                         dtype=torch.float16, timers=timers,
                         fp16_config=_fp16_cfg(static_loss_scale, dynamic_loss_scale, dynamic_loss_args,
                                               initial_dynamic_scale))
        self.fused_adam_legacy = fused_adam_legacy
