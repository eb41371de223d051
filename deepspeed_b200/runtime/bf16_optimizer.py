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

    # ---- reference method names (``bf16_optimizer.py:37-590``) over the flat-unit state ----------------------------
    @torch.no_grad()
    def clear_hp_grads(self):
        """Zero the fp32 gradient accumulators."""
        for st in self.storages.values():
            if st.grad is not None:
                st.grad.zero_()
        for u in self.units:
            u.has_grad_data = False

    def clear_lp_grads(self, set_to_none=False):
        """Drop / zero the low-precision ``.grad`` of the module parameters (they are consumed by the hooks anyway)."""
        for group in self.bf16_groups:
            for p in group:
                if set_to_none or p.grad is None:
                    p.grad = None
                else:
                    p.grad.zero_()

    def update_hp_grads(self, clear_lp_grads=False):
        """Gradients move into the fp32 accumulators as they are produced (per-parameter hooks), so at this point there
        is nothing left to copy; kept for callers of the reference's two-phase API."""
        if clear_lp_grads:
            self.clear_lp_grads()

    def get_grads_for_reduction(self):
        """(dense gradient tensors, {expert group: tensors}) that a plain data-parallel all-reduce would cover."""
        dense = [u.full_grad for u in self.units if u.full_grad is not None]
        return dense, {}

    def get_grads_for_norm(self, for_clipping=False):
        """This rank's fp32 gradient shards (their norms combine to the global gradient norm)."""
        return [st.grad for st in self.storages.values() if st.grad is not None]
