"""``BF16_Optimizer`` (reference ``runtime/bf16_optimizer.py:37``): bf16 parameters with fp32 master weights and fp32
gradient accumulation, optimizer state partitioned over the data-parallel group. In this framework that is exactly
``ZeroShardedOptimizer(stage=1, dtype=bf16, grad_accum_dtype=fp32)``; the class keeps the reference's constructor."""
import torch

from .zero.sharded_optimizer import ZeroShardedOptimizer


class BF16_Optimizer(ZeroShardedOptimizer):

    def __init__(self, init_optimizer, param_names=None, mpu=None, clip_grad=0.0, norm_type=2,
                 allgather_bucket_size=5000000000, dp_process_group=None, timers=None, grad_acc_dtype=None,
                 graph_harvesting=False, immediate_grad_update=False, has_moe_layers=False, module=None):
        from .. import comm as dist
        world = dist.get_world_size(dp_process_group) if dist.is_initialized() else 1
        rank = dist.get_rank(dp_process_group) if dist.is_initialized() else 0
        assert module is not None, "BF16_Optimizer needs the module whose parameters it manages (module=...)"
        super().__init__(module, init_optimizer, stage=1, dp_group=dp_process_group, dp_world=world, dp_rank=rank,
                         mp_group=(mpu.get_model_parallel_group() if mpu is not None else None), clip_grad=clip_grad,
                         dtype=torch.bfloat16, grad_accum_dtype=grad_acc_dtype or torch.float32, timers=timers)
        self.bf16_groups = [list(g) for g in self._orig_param_groups]
        self.immediate_grad_update = immediate_grad_update
