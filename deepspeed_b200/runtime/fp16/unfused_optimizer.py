"""``FP16_UnfusedOptimizer`` (reference ``runtime/fp16/unfused_optimizer.py:24``): the per-tensor (LAMB-compatible)
variant; same wrapper here because the inner optimizer decides whether the update is fused."""
from .fused_optimizer import FP16_Optimizer


class FP16_UnfusedOptimizer(FP16_Optimizer):

    def __init__(self, init_optimizer, deepspeed=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0, fused_lamb_legacy=False, module=None):
        super().__init__(init_optimizer, deepspeed=deepspeed, static_loss_scale=static_loss_scale,
                         dynamic_loss_scale=dynamic_loss_scale, dynamic_loss_args=dynamic_loss_args, verbose=verbose,
                         mpu=mpu, clip_grad=clip_grad, module=module)
        self.fused_lamb_legacy = fused_lamb_legacy
