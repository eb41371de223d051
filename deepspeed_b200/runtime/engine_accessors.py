"""The remaining ~90 configuration accessors and small helpers of the reference engine
(``runtime/engine.py:989-1482`` and scattered helpers), kept out of ``engine.py`` so the orchestration code stays
readable. Every accessor reads the typed config; a section this framework does not carry answers with the reference's
default instead of raising."""
import os

import torch

from .. import comm as dist
from ..utils import groups


def _get(obj, path, default=None):
    for part in path.split("."):
        if obj is None:
            return default
        if isinstance(obj, dict):
            if part not in obj:
                return default
            obj = obj[part]
        else:
            if not hasattr(obj, part):
                return default
            obj = getattr(obj, part)
    return obj


def _accessor(path, default=None):

    def fn(self):
        return _get(self._config, path, default)

    fn.__doc__ = f"ds_config value ``{path}`` (reference default {default!r})."
    return fn


class EngineAccessorsMixin:
    # ---- plain config reads -------------------------------------------------------------------
    aio_config = _accessor("aio_config", {})
    amp_params = _accessor("amp_params", None)
    autotuning_end_profile_step = _accessor("autotuning_config.end_profile_step", 5)
    autotuning_start_profile_step = _accessor("autotuning_config.start_profile_step", 3)
    autotuning_metric = _accessor("autotuning_config.metric", "throughput")
    bf16_master_weights_and_gradients = _accessor("bfloat16_config.bf16_master_weights_and_grads", False)
    bf16_optimizer_states = _accessor("bfloat16_config.bf16_optimizer_states", False)
    curriculum_params_legacy = _accessor("curriculum_params_legacy", None)
    dataloader_drop_last = _accessor("dataloader_drop_last", False)
    eigenvalue_gas_boundary_resolution = _accessor("eigenvalue_config.gas_boundary_resolution", 1)
    eigenvalue_layer_name = _accessor("eigenvalue_config.layer_name", "bert.encoder.layer")
    eigenvalue_layer_num = _accessor("eigenvalue_config.layer_num", 0)
    eigenvalue_max_iter = _accessor("eigenvalue_config.max_iter", 100)
    eigenvalue_stability = _accessor("eigenvalue_config.stability", 1e-6)
    eigenvalue_tol = _accessor("eigenvalue_config.tol", 1e-2)
    eigenvalue_verbose = _accessor("eigenvalue_config.verbose", False)
    flops_profiler_module_depth = _accessor("flops_profiler_config.module_depth", -1)
    flops_profiler_output_file = _accessor("flops_profiler_config.output_file", None)
    flops_profiler_recompute_fwd_factor = _accessor("flops_profiler_config.recompute_fwd_factor", 0.0)
    flops_profiler_top_modules = _accessor("flops_profiler_config.top_modules", 1)
    fp16_auto_cast = _accessor("float16_config.auto_cast", False)
    fp16_master_weights_and_gradients = _accessor("float16_config.fp16_master_weights_and_grads", False)
    graph_harvesting = _accessor("graph_harvesting", False)
    log_level = _accessor("log_level", "info")
    managed_gradient_accumulation = _accessor("managed_gradient_accumulation", True)
    mics_shard_size = _accessor("mics_shard_size", -1)
    optimizer_legacy_fusion = _accessor("optimizer_legacy_fusion", False)
    pld_params = _accessor("pld_params", False)
    swap_tensor_config = _accessor("aio_config", None)  # this framework keeps the swapper settings in the aio section
    tensor_parallel_config = _accessor("tensor_parallel_config", None)
    torch_autocast_enabled = _accessor("torch_autocast_config.enabled", False)
    use_node_local_storage = _accessor("use_node_local_storage", False)
    zenflow_config = _accessor("zero_config.zenflow", None)
    zero_allgather_partitions = _accessor("zero_config.allgather_partitions", True)
    zero_allgather_sequential = _accessor("zero_config.allgather_sequential", False)
    zero_allow_untested_optimizer = _accessor("zero_allow_untested_optimizer", False)
    zero_force_ds_cpu_optimizer = _accessor("zero_force_ds_cpu_optimizer", True)
    zero_ignore_unused_parameters = _accessor("zero_config.ignore_unused_parameters", True)
    zero_legacy_stage1 = _accessor("zero_config.legacy_stage1", False)
    zero_log_trace_cache_warnings = _accessor("zero_config.log_trace_cache_warnings", False)
    zero_model_persistence_threshold = _accessor("zero_config.model_persistence_threshold", int(1e14))
    zero_module_granularity_threshold = _accessor("zero_config.module_granularity_threshold", 0)
    zero_multi_rank_bucket_allreduce = _accessor("zero_config.use_multi_rank_bucket_allreduce", True)
    zero_offload_optimizer = _accessor("zero_config.offload_optimizer", None)
    zero_quantized_nontrainable_weights = _accessor("zero_config.zero_quantized_nontrainable_weights", False)
    zero_round_robin_gradients = _accessor("zero_config.round_robin_gradients", False)
    zero_save_muon_momentum_buffer_in_memory = _accessor("zero_config.save_muon_momentum_buffer_in_memory", False)
    zeropp_loco_param = _accessor("zero_config.zeropp_loco_param", None)
    is_sanity_checks_enabled = _accessor("zero_config.enable_sanity_checks", False)
    zero_grad_hooks = _accessor("zero_config.grad_hooks", True)

    # ---- derived values -----------------------------------------------------------------------
    def torch_autocast_dtype(self):
        name = _get(self._config, "torch_autocast_config.dtype", None)
        if name is None or isinstance(name, torch.dtype):
            return name
        return {"bfloat16": torch.bfloat16, "bf16": torch.bfloat16, "float16": torch.float16, "fp16": torch.float16,
                "half": torch.float16, "float32": torch.float32, "fp32": torch.float32}.get(
                    str(name).replace("torch.", ""), None)

    def autotp_size(self):
        return _get(self._config, "tensor_parallel_config.autotp_size", 0)

    def autotuning_metric_path(self):
        return _get(self._config, "autotuning_config.metric_path", None) or \
            os.path.join(os.getcwd(), "autotuning_metric.json")

    def autotuning_model_info_path(self):
        return _get(self._config, "autotuning_config.model_info_path", None) or \
            os.path.join(os.getcwd(), "autotuning_model_info.json")

    def autotuning_profile_model_info(self):
        info = _get(self._config, "autotuning_config.model_info", None)
        return bool(self.autotuning_enabled() and info and info.get("profile", False))

    def checkpoint_serialization_enabled(self):
        return bool(_get(self._config, "checkpoint_config.checkpoint_serialization", True))

    def checkpoint_writer_enabled(self):
        return _get(self._config, "checkpoint_config.writer", None) is not None

    def cpuadam_cores_perc(self):
        return _get(self._config, "zero_config.offload_optimizer.cpuadam_cores_perc", 0.9)

    def super_offload(self):
        return bool(_get(self._config, "zero_config.offload_optimizer.super_offload", False))

    def zero_partial_offload(self):
        return _get(self._config, "zero_config.offload_optimizer.ratio", 1.0)

    def zero_use_cpu_optimizer(self):
        dev = _get(self._config, "zero_config.offload_optimizer.device", None)
        return dev is not None and str(getattr(dev, "value", dev)) in ("cpu", "nvme")

    def zero_nvme_offload_optimizer(self):
        return bool(getattr(self.optimizer, "swap_optimizer", False) or getattr(self.optimizer, "offload_nvme", False))

    def dynamic_loss_scale_args(self):
        fc = _get(self._config, "float16_config", None)
        if fc is not None and hasattr(fc, "dynamic_loss_scale_args"):
            return fc.dynamic_loss_scale_args()
        return _get(self._config, "dynamic_loss_scale_args", None)

    def initial_dynamic_scale(self):
        fc = _get(self._config, "float16_config", None)
        if fc is not None and hasattr(fc, "initial_dynamic_scale"):
            return fc.initial_dynamic_scale()
        return 2**_get(self._config, "float16_config.initial_scale_power", 16)

    def flops_profiler_detailed(self):
        if self.autotuning_enabled():
            return False
        return _get(self._config, "flops_profiler_config.detailed", True)

    def quantize_training(self):
        """The MoQ tuple of the reference (``engine.py:1120``): in-forward flag, enabled, groups, fp16 mixed, change
        ratio, type, rounding, verbose, kernel."""
        sp = _get(self._config, "compression_config.weight_quantization.shared_parameters", {}) or {}
        g = sp.get
        return (g("quantize_weight_in_forward", False), g("enabled", False), g("quantize_groups", 1),
                g("fp16_mixed_quantize", {}).get("enabled", False) if isinstance(g("fp16_mixed_quantize", {}), dict)
                else False, g("fp16_mixed_quantize", {}).get("quantize_change_ratio", 0.001)
                if isinstance(g("fp16_mixed_quantize", {}), dict) else 0.001, g("quantization_type", "symmetric"),
                g("rounding", "nearest"), g("quantize_verbose", False), g("quantizer_kernel", False))

    def torch_autocast_lower_precision_safe_modules(self):
        names = _get(self._config, "torch_autocast_config.lower_precision_safe_modules", None)
        if names is None:
            from .torch_autocast import get_default_autocast_lower_precision_modules
            return get_default_autocast_lower_precision_modules()
        return names

    def is_deepcompile_enabled(self):
        return bool(_get(self._config, "compile_config.deepcompile", False))

    def is_deepcompile_active(self):
        return getattr(self, "_dc_driver", None) is not None

    def compile_zero_optimization_stage(self):
        passes = _get(self._config, "compile_config.passes", None) or []
        return "z1" in passes or "z3" in passes

    def compile_autosp(self):
        return "autosp" in (_get(self._config, "compile_config.passes", None) or [])

    def compile_autotp(self):
        return "autotp" in (_get(self._config, "compile_config.passes", None) or [])

    def uses_parallelization_pass_only(self):
        return self.compile_autosp() or self.compile_autotp()

    def is_elastic_model_parallel_supported(self):
        if self.elasticity_enabled():
            mp = _get(self._config, "elastic_model_parallel_size", 1) or 1
            return _get(self._config, "num_gpus_per_node", 1) % mp == 0
        return False

    def is_first_weights_partition_group(self):
        ret = self.mics_shard_size() < 0 and self.zero_optimization_partition_weights()
        if self.mics_shard_size() > 0 and self.global_rank < self.mics_shard_size():
            ret = True
        return ret

    # ---- small helpers ------------------------------------------------------------------------
    @staticmethod
    def is_map_style_dataset(obj):
        return hasattr(obj, "__getitem__") and hasattr(obj, "__len__")

    @staticmethod
    def is_iterable_style_dataset(obj):
        return isinstance(obj, torch.utils.data.IterableDataset)

    def get_data_parallel_rank(self):
        return groups.get_data_parallel_rank()

    def get_model_parallel_rank(self):
        return groups.get_model_parallel_rank()

    def get_tensor_parallel_rank(self):
        return groups.get_tensor_model_parallel_rank()

    def get_pld_theta(self):
        return self.progressive_layer_drop.get_theta() if self.progressive_layer_drop else None

    def active_timers(self):
        return self.engine_timers.micro_timers + self.engine_timers.global_timers

    def get_wall_clock_timers(self):
        """Snapshot ``{timer name: elapsed ms}`` of the engine's wall-clock timers (reference ``engine.py:3541``)."""
        out = {}
        for name in self.active_timers():
            t = self.timers.timers.get(name) if hasattr(self.timers, "timers") else None
            if t is not None:
                out[name] = t.elapsed(reset=False) if hasattr(t, "elapsed") else 0.0
        return out

    def all_gather_scalar(self, value, dp_group):
        tensor_list = [value.new_zeros(value.size()) for _ in range(dist.get_world_size(group=dp_group))]
        dist.all_gather(tensor_list, value, group=dp_group)
        return tensor_list

    def clip_fp32_gradients(self):
        from .utils import clip_grad_norm_
        clip_grad_norm_(parameters=self.module.parameters(), max_norm=self.gradient_clipping(), mpu=self.mpu)

    def dump_state(self):
        """Reference ``engine.py:1573``: one line per engine attribute on rank 0."""
        from ..utils.logging import log_dist
        for k, v in sorted(vars(self).items()):
            if not k.startswith("_") and not isinstance(v, (torch.nn.Module, torch.Tensor)):
                log_dist(f"  {k} = {v!r:.200}", [0])

    def empty_partition_cache(self):
        """Release the ZeRO-3 parameters that are still gathered (reference ``engine.py:5698``)."""
        co = getattr(self.optimizer, "coordinator", None)
        if co is not None:
            co.release_all()

    # ---- plain data-parallel reduction helpers (reference engine.py:3658-3870) ------------------
    def allreduce_bucket(self, bucket, dp_group, dp_world_size=None):
        """Flatten ``bucket``, all-reduce it in the communication dtype with the configured averaging
        (pre-/post-scaling by ``gradient_predivide_factor``) and return the flat result."""
        from torch._utils import _flatten_dense_tensors
        tensor = _flatten_dense_tensors(bucket)
        work = tensor if self.communication_data_type == tensor.dtype else tensor.to(self.communication_data_type)
        if dp_world_size is None:
            dp_world_size = dist.get_world_size(group=dp_group)
        if not self.gradient_average:
            dist.all_reduce(work, group=dp_group)
        elif self.postscale_gradients():
            pre = self.gradient_predivide_factor()
            if pre != 1.0:
                work.mul_(1.0 / pre)
            dist.all_reduce(work, group=dp_group)
            if pre != dp_world_size:
                work.mul_(pre / dp_world_size)
        else:
            work.mul_(1.0 / dp_world_size)
            dist.all_reduce(work, group=dp_group)
        if work is not tensor:
            tensor.copy_(work)
        return tensor

    def allreduce_and_copy(self, small_bucket, dp_group, dp_world_size=None):
        from torch._utils import _unflatten_dense_tensors
        reduced = self.allreduce_bucket(small_bucket, dp_group, dp_world_size)
        for buf, synced in zip(small_bucket, _unflatten_dense_tensors(reduced, small_bucket)):
            buf.copy_(synced)

    def allreduce_no_retain(self, bucket, dp_group, numel_per_bucket=500000000, dp_world_size=None):
        small, numel = [], 0
        for t in bucket:
            small.append(t)
            numel += t.numel()
            if numel > numel_per_bucket:
                self.allreduce_and_copy(small, dp_group, dp_world_size)
                small, numel = [], 0
        if small:
            self.allreduce_and_copy(small, dp_group, dp_world_size)

    def buffered_allreduce_fallback(self, grads=None, elements_per_buffer=500000000):
        """All-reduce ``param.grad`` of the wrapped module (or ``grads``) over the data-parallel group in dtype
        buckets; expert parameters go over their expert-data-parallel groups."""
        if grads is None:
            dense, expert = [], {}
            for _, p in self.module.named_parameters():
                if not p.requires_grad or p.numel() == 0:
                    continue
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
                if getattr(p, "allreduce", True) is False and hasattr(p, "group_name"):
                    expert.setdefault(p.group_name, []).append(p.grad.data)
                else:
                    dense.append(p.grad.data)
        else:
            assert not self.has_moe_layers, "attempting to reduce grads in unsupported way w.r.t. MoE"
            dense, expert = list(grads), {}
        by_dtype = {}
        for g in dense:
            by_dtype.setdefault(g.dtype, []).append(g)
        dp_group = groups.get_data_parallel_group()
        for bucket in by_dtype.values():
            self.allreduce_no_retain(bucket, dp_group, elements_per_buffer)
        for name, gs in expert.items():
            eg = groups._get_expert_data_parallel_group(name)
            by_dtype = {}
            for g in gs:
                by_dtype.setdefault(g.dtype, []).append(g)
            for bucket in by_dtype.values():
                self.allreduce_no_retain(bucket, eg, elements_per_buffer)

    def sparse_all_gather(self, value, dp_group):
        my_size = torch.tensor([value.size(0)], dtype=torch.long, device=value.device)
        all_sizes = self.all_gather_scalar(my_size, dp_group)
        max_size = int(torch.cat(all_sizes).max())
        fill = max_size - value.size(0)
        assert value.dim() in (1, 2)
        if fill > 0:
            pad_shape = (fill, ) if value.dim() == 1 else (fill, value.size(1))
            value = torch.cat([value, value.new_zeros(pad_shape)])
        shape = (max_size, ) if value.dim() == 1 else (max_size, value.size(1))
        tensor_list = [value.new_empty(shape) for _ in range(dist.get_world_size(group=dp_group))]
        dist.all_gather(tensor_list, value.contiguous(), group=dp_group)
        return [t[:int(all_sizes[i][0])] for i, t in enumerate(tensor_list)]

    def sparse_allreduce(self, sparse, dp_group, dp_world_size=None):
        """``SparseTensor`` (indices + values of an embedding gradient): scale, all-gather both parts."""
        orig = sparse.values.dtype
        indices, values = sparse.indices, sparse.values
        if self.communication_data_type != values.dtype:
            if self.communication_data_type in (torch.float16, torch.bfloat16):
                indices = indices.to(torch.int32)
            values = values.to(self.communication_data_type)
        if dp_world_size is None:
            dp_world_size = dist.get_world_size(group=dp_group)
        if self.gradient_average:
            if self.postscale_gradients():
                values = values * (self.gradient_predivide_factor() / dp_world_size)
            else:
                values = values * (1.0 / dp_world_size)
        sparse.indices = torch.cat(self.sparse_all_gather(indices, dp_group)).to(torch.long)
        sparse.values = torch.cat(self.sparse_all_gather(values, dp_group)).to(orig)
        return sparse

    def sparse_allreduce_bucket(self, bucket, dp_group, dp_world_size=None):
        return [self.sparse_allreduce(sp, dp_group, dp_world_size) for sp in bucket]

    def sparse_allreduce_no_retain(self, bucket, dp_group, dp_world_size=None):
        for t in self.sparse_allreduce_bucket(bucket, dp_group, dp_world_size):
            if getattr(t, "is_sparse", False):
                t.orig_dense_tensor.data = t.to_coo_tensor()
            else:
                t.orig_dense_tensor.copy_(t.to_dense())

    # ---- misc ---------------------------------------------------------------------------------
    def print_forward_breakdown(self, fwd_time):
        from ..utils.logging import log_dist
        gate = sum(getattr(g, "gate_time", 0.0) for g in self.gate_modules)
        moe = sum(getattr(m, "time_moe", 0.0) for m in self.moe_layers)
        a1 = sum(getattr(m, "time_falltoall", 0.0) for m in self.moe_layers)
        a2 = sum(getattr(m, "time_salltoall", 0.0) for m in self.moe_layers)
        log_dist(f"time (ms) | fwd: {fwd_time:.2f} (fwd_moe: {moe:.2f}, 1st_a2a: {a1:.2f}, 2nd_a2a: {a2:.2f}, "
                 f"top_k: {gate:.2f})", ranks=[0])

    def save_fp16_model(self, save_dir, save_filename="pytorch_model.bin"):
        """Deprecated name of :meth:`save_16bit_model` (reference ``engine.py:5562``)."""
        return self.save_16bit_model(save_dir, save_filename)

    def get_compile_time(self):
        """Seconds spent per DeepCompile stage (profiling result gathering, planning)."""
        drv = getattr(self, "_dc_driver", None)
        return dict(getattr(drv, "pass_times", {})) if drv is not None else {}

    def register_compile_pass(self, pass_name, pass_fn, contract=None):
        """Register a custom pass for the unit plan: ``pass_fn(profile, plan) -> plan | None`` runs after the built-in
        planner when ``pass_name`` appears in the ``engine.compile(schedule=...)`` pass list (reference
        ``engine.py:5798``; there the pass rewrites an FX graph, here the :class:`compile.planner.UnitPlan`)."""
        from ..compile.backend import register_compile_pass
        register_compile_pass(pass_name, pass_fn, contract)

    @staticmethod
    def load_moe_state_dict(checkpoint_path, tag, state_dict, old_moe_load, model=None, mpu=None, num_experts=1,
                            checkpoint_engine=None):
        """Merge the per-expert checkpoint files of this rank's experts into ``state_dict`` (reference
        ``engine.py:3994``)."""
        from ..moe.checkpoint import load_expert_files
        return load_expert_files(checkpoint_path, tag, state_dict, model=model, mpu=mpu, num_experts=num_experts,
                                 old_moe_load=old_moe_load)
