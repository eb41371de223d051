"""Reference-named public methods of the ZeRO optimizers on :class:`ZeroShardedOptimizer`
(``runtime/zero/stage3.py``, ``stage_1_and_2.py``): learning-rate access, loss-scale override, overflow checks, the
per-parameter fp32 / gradient accessors behind ``utils.safe_get_*``, partition-cache control. Thin adapters over the
flat-unit bookkeeping; kept apart from the hot-path code."""
import torch

from ... import comm as dist
from ...utils.logging import logger


class ZeroCompatAPI:

    # ---- learning rate / loss scale ------------------------------------------------------------------
    def set_lr(self, lr):
        for g in self.optimizer.param_groups:
            g["lr"] = lr

    def get_lr(self):
        return self.optimizer.param_groups[0]["lr"]

    def override_loss_scale(self, loss_scale):
        if loss_scale != self.external_loss_scale:
            logger.info(f"[deepspeed] setting loss scale from {self.external_loss_scale} -> {loss_scale}")
        self.custom_loss_scaler = True
        self.external_loss_scale = loss_scale

    def _get_state(self):
        return self.optimizer.state

    def _set_state(self, value):
        self.optimizer.state = value

    state = property(_get_state, _set_state)

    # ---- overflow ------------------------------------------------------------------------------------
    @torch.no_grad()
    def has_overflow(self, partition_gradients=True):
        """inf / nan anywhere in the reduced gradient shards, agreed over the data-parallel (and model-parallel) ranks."""
        bad = False
        for st in self.storages.values():
            g = st.grad
            if g is not None and g.numel():
                bad = bad or not bool(torch.isfinite(g.float().sum()))
        flag = torch.tensor([1.0 if bad else 0.0], device=next(iter(self.storages.values())).device
                            if self.storages else "cpu")
        if dist.is_initialized():
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
        return bool(flag.item() > 0)

    def check_overflow(self, partition_gradients=True):
        self.overflow = self.has_overflow(partition_gradients)

    def has_overflow_serial(self, params, is_grad_list=False):
        for p in params:
            g = p if is_grad_list else p.grad
            if g is not None and not bool(torch.isfinite(g.float().sum())):
                return True
        return False

    @torch.no_grad()
    def unscale_and_clip_grads(self, sub_group_id=None, total_norm=None):
        """Divide the gradient shards by the loss scale combined with the clipping coefficient (``total_norm`` is the
        norm of the SCALED gradients)."""
        scale = float(self.cur_scale)
        combined = scale
        if self.clip_grad > 0.0 and total_norm is not None:
            clip = max(1.0, (float(total_norm) / scale + 1e-6) / self.clip_grad)
            combined = clip * scale
        items = list(self.storages.values())
        if sub_group_id is not None:
            items = items[sub_group_id:sub_group_id + 1]
        for st in items:
            if st.grad is not None:
                st.grad.mul_(1.0 / combined)

    # ---- parameters / gradients by parameter ------------------------------------------------------------
    def get_param_id(self, param):
        u, i = self._unit_of_param[id(param)]
        return (u.uid, i)

    def get_fp32_grad_for_param(self, param):
        """Full fp32 gradient of ``param`` (collective)."""
        return self.get_full_grad(param)

    def set_fp32_grad_for_param(self, value, param):
        from ...utils.tensor_fragment import safe_set_full_grad
        safe_set_full_grad(param, value)

    def get_local_fp32_param(self, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        if optim_state_key is None:
            return tf.safe_get_local_fp32_param(param)
        return tf.safe_get_local_optimizer_state(param, optim_state_key)

    def set_local_hp_param(self, value, param, optim_state_key=None):
        from ...utils import tensor_fragment as tf
        if optim_state_key is None:
            tf.safe_set_local_fp32_param(param, value)
        else:
            tf.safe_set_local_optimizer_state(param, value, optim_state_key)

    def get_local_fp32_grad_for_param(self, param):
        from ...utils.tensor_fragment import safe_get_local_grad
        return safe_get_local_grad(param)

    def set_local_grad_for_param(self, value, param):
        from ...utils.tensor_fragment import safe_set_local_grad
        safe_set_local_grad(param, value)

    def update_fp32_grad_for_param_vectorized(self, update_func, param_list):
        from ...utils.tensor_fragment import safe_update_full_grad_vectorized
        safe_update_full_grad_vectorized(param_list, update_func)

    def get_hp_param_device(self, param, optim_state_key=None):
        if not param.requires_grad:
            return None
        u, _ = self._unit_of_param[id(param)]
        if optim_state_key is None:
            return u.master_shard.device
        s = self._state_shard(u, optim_state_key)
        return None if s is None else s.device

    # ---- fp32 <-> low precision --------------------------------------------------------------------------
    def refresh_fp32_params(self):
        """Rebuild the fp32 master shards from the current low-precision parameters (after loading module weights)."""
        co = self.coordinator
        if co is not None:
            co.gather_all()
        self.refresh_lp_from_params()
        if co is not None:
            co.release_all()

    def update_lp_params(self):
        """Copy the master shards into the low-precision parameters (and re-gather what is replicated)."""
        self._join_step()
        self._lp_written_by_step = False  # the masters were changed outside step(): the copy must really happen
        self._lp_shards_will_be_written()
        self._sync_lp_from_master()
        self._post_step_gather()

    def initialize_optimizer_states(self):
        """Materialise the inner optimizer's state with a zero-gradient step that leaves the parameters unchanged
        (the reference does this at construction so that checkpoints of step 0 already contain the state)."""
        saved = [st.master.clone() for st in self.storages.values()]
        for st in self.storages.values():
            if st.grad is not None:
                st.grad.zero_()
        lrs = [g["lr"] for g in self.optimizer.param_groups]
        for g in self.optimizer.param_groups:
            g["lr"] = 0.0
        try:
            self.optimizer.step()
        finally:
            for g, lr in zip(self.optimizer.param_groups, lrs):
                g["lr"] = lr
            for st, m in zip(self.storages.values(), saved):
                st.master.copy_(m)
            for state in self.optimizer.state.values():
                if "step" in state:
                    state["step"] = state["step"] * 0 if torch.is_tensor(state["step"]) else 0

    # ---- accumulation boundary / caches ------------------------------------------------------------------
    def set_gradient_accumulation_boundary(self, is_boundary):
        self.is_gradient_accumulation_boundary = is_boundary

    def finalize_gradient_accumulation_boundary(self):
        """Caller-owned accumulation (``managed_gradient_accumulation: false``): gradients of stages 2/3 are reduced at
        every backward already, so marking the boundary is all that is left."""
        self.set_gradient_accumulation_boundary(True)

    def empty_partition_cache(self):
        if self.coordinator is not None:
            self.coordinator.release_all()

    def invalidate_secondary_tensor(self):
        """hpZ: drop the intra-node secondary shards (stale after an update)."""
        for u in self.units:
            u.sec_shard = None

    def print_rank_0(self, message):
        if not dist.is_initialized() or dist.get_rank() == 0:
            logger.info(message)

    def get_lean_optimizer_state(self):
        """Inner optimizer state of this rank without padding: ``{key: [tensor per param group]}`` (fp32 masters under
        ``"fp32"``)."""
        from ...checkpoint.layout import ReferenceLayout
        lay = ReferenceLayout(self)
        out = {"fp32": lay.export(lambda u: u.master_shard)}
        keys = sorted({k for st in self.optimizer.state.values() for k, v in st.items()
                       if torch.is_tensor(v) and v.numel() > 1})
        for k in keys:
            out[k] = lay.export(lambda u, k=k: self._state_shard(u, k))
        return out
