"""Dictionary readers of the ds_config: ``get_<field>(param_dict)``.

Reference: the module-level getters of ``runtime/config.py`` (and ``runtime/precision_config.py``,
``runtime/data_pipeline/config.py``, ``monitor/config.py``, ``runtime/tensor_parallel/config.py``). The engine of this
framework reads the typed :class:`DeepSpeedConfig`; these functions serve code that inspects a raw config dictionary
(launch scripts, autotuners, integrations) with the reference's defaults."""
import copy


def _p(d, key, default=None):
    return d.get(key, default) if isinstance(d, dict) else default


def _sub(d, key):
    v = _p(d, key, {})
    return v if isinstance(v, dict) else {}


# ---- batch / schedule -------------------------------------------------------------------------------------------
def get_train_batch_size(param_dict):
    return _p(param_dict, "train_batch_size")


def get_train_micro_batch_size_per_gpu(param_dict):
    return _p(param_dict, "train_micro_batch_size_per_gpu")


def get_gradient_accumulation_steps(param_dict):
    return _p(param_dict, "gradient_accumulation_steps")


def get_managed_gradient_accumulation(param_dict):
    return _p(param_dict, "managed_gradient_accumulation", True)


def get_steps_per_print(param_dict):
    return _p(param_dict, "steps_per_print", None)


def get_dump_state(param_dict):
    return _p(param_dict, "dump_state", False)


def get_wall_clock_breakdown(param_dict):
    return _p(param_dict, "wall_clock_breakdown", False)


def get_memory_breakdown(param_dict):
    return _p(param_dict, "memory_breakdown", False)


def get_log_level(param_dict):
    return _p(param_dict, "log_level", None)


def get_dataloader_drop_last(param_dict):
    return _p(param_dict, "dataloader_drop_last", False)


# ---- gradients / communication ------------------------------------------------------------------------------------
def get_gradient_clipping(param_dict):
    return _p(param_dict, "gradient_clipping", 0.0)


def get_prescale_gradients(param_dict):
    return _p(param_dict, "prescale_gradients", False)


def get_gradient_predivide_factor(param_dict):
    return _p(param_dict, "gradient_predivide_factor", 1.0)


def get_gradient_allreduce_op(param_dict):
    op = str(_p(param_dict, "gradient_allreduce_op", "avg")).lower()
    if op not in ("avg", "sum"):
        raise ValueError(f"gradient_allreduce_op must be 'avg' or 'sum', got {op!r}")
    return op


def get_sparse_gradients_enabled(param_dict):
    return _p(param_dict, "sparse_gradients", False)


def get_disable_allgather(param_dict):
    return _p(param_dict, "disable_allgather", False)


def get_communication_data_type(param_dict, key="communication_data_type", default=None):
    import torch
    v = _p(param_dict, key, default)
    if v is None:
        return None
    table = {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16, "bfp16": torch.bfloat16}
    if v not in table:
        raise ValueError(f"Invalid communication_data_type. Supported data types: ['fp16', 'bf16', 'fp32']. Got: {v}")
    return table[v]


def get_graph_harvesting(param_dict):
    return _p(param_dict, "graph_harvesting", False)


# ---- precision ---------------------------------------------------------------------------------------------------
def get_float16_config(param_dict):
    from .sections import DeepSpeedFP16Config as FP16Config
    return FP16Config(**_sub(param_dict, "fp16"))


def get_bfloat16_config(param_dict):
    from .sections import DeepSpeedBF16Config as BF16Config
    d = _sub(param_dict, "bf16") or _sub(param_dict, "bfloat16")
    return BF16Config(**d)


def get_amp_enabled(param_dict):
    return _p(_sub(param_dict, "amp"), "enabled", False)


def get_amp_params(param_dict):
    amp = copy.copy(_sub(param_dict, "amp"))
    amp.pop("enabled", None)
    return amp if "amp" in (param_dict or {}) else False


def get_torch_autocast_enabled(param_dict):
    return _p(_sub(param_dict, "torch_autocast"), "enabled", False)


def get_torch_autocast_dtype(param_dict):
    import torch
    name = _p(_sub(param_dict, "torch_autocast"), "dtype", None)
    if name is None:
        return None
    return {"bfloat16": torch.bfloat16, "bf16": torch.bfloat16, "float16": torch.float16, "fp16": torch.float16,
            "half": torch.float16, "float32": torch.float32}[str(name).replace("torch.", "")]


def get_lower_precision_safe_modules(param_dict):
    return _p(_sub(param_dict, "torch_autocast"), "lower_precision_safe_modules", None)


def get_data_types_params(param_dict):
    return _sub(param_dict, "data_types")


# ---- optimizer / scheduler ---------------------------------------------------------------------------------------
def get_optimizer_name(param_dict):
    name = _p(_sub(param_dict, "optimizer"), "type", None)
    return name.lower() if isinstance(name, str) and name.lower() in ("adam", "adamw", "lamb", "onebitadam", "onebitlamb",
                                                                        "zerooneadam", "adagrad", "lion", "muon",
                                                                        "sgd") else name


def get_optimizer_params(param_dict):
    return _p(_sub(param_dict, "optimizer"), "params", None) if get_optimizer_name(param_dict) is not None else None


def get_optimizer_gradient_clipping(param_dict):
    params = get_optimizer_params(param_dict)
    return params.get("max_grad_norm") if isinstance(params, dict) else None


def get_optimizer_legacy_fusion(param_dict):
    return _p(_sub(param_dict, "optimizer"), "legacy_fusion", False)


def get_zero_allow_untested_optimizer(param_dict):
    return _p(param_dict, "zero_allow_untested_optimizer", False)


def get_zero_force_ds_cpu_optimizer(param_dict):
    return _p(param_dict, "zero_force_ds_cpu_optimizer", True)


def get_scheduler_name(param_dict):
    return _p(_sub(param_dict, "scheduler"), "type", None)


def get_scheduler_params(param_dict):
    return _p(_sub(param_dict, "scheduler"), "params", None) if get_scheduler_name(param_dict) is not None else None


# ---- sections ----------------------------------------------------------------------------------------------------
def get_pipeline_config(param_dict):
    cfg = {"stages": "auto", "partition": "best", "seed_layers": False, "activation_checkpoint_interval": 0,
           "pipe_partitioned": True, "grad_partitioned": True}
    cfg.update(_sub(param_dict, "pipeline"))
    return cfg


def get_hybrid_engine_config(param_dict):
    from .sections import HybridEngineConfig
    return HybridEngineConfig(**_sub(param_dict, "hybrid_engine"))


def get_expert_parallel_config(param_dict):
    from ..parallel.auto_ep_presets import parse_autoep_config
    return parse_autoep_config(_sub(param_dict, "expert_parallel"))


def get_expert_data_topo_config(param_dict):
    return _p(param_dict, "use_data_before_expert_parallelism", False)


def get_tensor_parallel_config(param_dict):
    from .sections import TPTrainingConfig
    return TPTrainingConfig(**_sub(param_dict, "tensor_parallel"))


def get_monitor_config(param_dict):
    from .sections import DeepSpeedMonitorConfig
    keys = ("tensorboard", "wandb", "csv_monitor", "comet")
    return DeepSpeedMonitorConfig(**{k: param_dict[k] for k in keys if k in (param_dict or {})})


def get_checkpoint_params(param_dict):
    return _sub(param_dict, "checkpoint")


def get_checkpoint_tag_validation_mode(checkpoint_params):
    mode = str(_p(checkpoint_params, "tag_validation", "WARN")).upper()
    if mode not in ("WARN", "IGNORE", "FAIL"):
        raise ValueError(f"Checkpoint config contains invalid tag_validation value of {mode}, expecting one of "
                         "['WARN', 'IGNORE', 'FAIL']")
    return mode


def get_checkpoint_parallel_write_pipeline(checkpoint_params):
    pw = _sub(checkpoint_params, "parallel_write")
    v = _p(pw, "pipeline_stage", False)
    if v not in (True, False):
        raise ValueError(f"checkpoint::parallel_write::pipeline_stage value of '{v}' is invalid, expecting: true or "
                         "false")
    return v


# ---- sparse attention ----------------------------------------------------------------------------------------------
_SPARSE_COMMON = (("block", 16), ("different_layout_per_head", False))
_SPARSE_FIELDS = {
    "dense": (("block", 16), ),
    "fixed": _SPARSE_COMMON + (("num_local_blocks", 4), ("num_global_blocks", 1), ("attention", "bidirectional"),
                               ("horizontal_global_attention", False), ("num_different_global_patterns", 1)),
    "variable": _SPARSE_COMMON + (("num_random_blocks", 0), ("local_window_blocks", [4]), ("global_block_indices", [0]),
                                  ("global_block_end_indices", None), ("attention", "bidirectional"),
                                  ("horizontal_global_attention", False)),
    "bigbird": _SPARSE_COMMON + (("num_random_blocks", 1), ("num_sliding_window_blocks", 3), ("num_global_blocks", 1)),
    "bslongformer": _SPARSE_COMMON + (("num_sliding_window_blocks", 3), ("global_block_indices", [0]),
                                      ("global_block_end_indices", None)),
}


def get_sparse_attention_mode(param_dict):
    return _p(param_dict, "mode", "fixed")


def get_sparse_attention_type(param_dict):
    return _p(param_dict, "attention", "bidirectional")


def _sparse(mode, sparsity):
    out = {"mode": mode}
    for key, default in _SPARSE_FIELDS[mode]:
        out[key] = _p(sparsity, key, copy.copy(default))
    return out


def get_sparse_dense_config(sparsity):
    return _sparse("dense", sparsity)


def get_sparse_fixed_config(sparsity):
    return _sparse("fixed", sparsity)


def get_sparse_variable_config(sparsity):
    return _sparse("variable", sparsity)


def get_sparse_bigbird_config(sparsity):
    return _sparse("bigbird", sparsity)


def get_sparse_bslongformer_config(sparsity):
    return _sparse("bslongformer", sparsity)


def get_sparse_attention(param_dict):
    if not isinstance(param_dict, dict) or "sparse_attention" not in param_dict:
        return None
    sparsity = param_dict["sparse_attention"]
    mode = get_sparse_attention_mode(sparsity)
    if mode not in _SPARSE_FIELDS:
        raise NotImplementedError(f"Given sparsity mode, {mode}, has not been implemented yet!")
    return _sparse(mode, sparsity)


# ---- eigenvalue / progressive layer drop -----------------------------------------------------------------------------
_EIGEN = (("enabled", False), ("verbose", False), ("max_iter", 100), ("tol", 1e-2), ("stability", 1e-6),
          ("gas_boundary_resolution", 1), ("layer_name", "bert.encoder.layer"), ("layer_num", 0))


def _eig(param_dict, key):
    return _p(_sub(param_dict, "eigenvalue"), key, dict(_EIGEN)[key])


def get_eigenvalue_enabled(param_dict):
    return _eig(param_dict, "enabled")


def get_eigenvalue_verbose(param_dict):
    return _eig(param_dict, "verbose")


def get_eigenvalue_max_iter(param_dict):
    return _eig(param_dict, "max_iter")


def get_eigenvalue_tol(param_dict):
    return _eig(param_dict, "tol")


def get_eigenvalue_stability(param_dict):
    return _eig(param_dict, "stability")


def get_eigenvalue_gas_boundary_resolution(param_dict):
    return _eig(param_dict, "gas_boundary_resolution")


def get_eigenvalue_layer_name(param_dict):
    return _eig(param_dict, "layer_name")


def get_eigenvalue_layer_num(param_dict):
    return _eig(param_dict, "layer_num")


def get_eigenvalue_config(param_dict):
    """(enabled, verbose, max_iter, tol, stability, gas_boundary_resolution, layer_name, layer_num)."""
    if get_quantize_enabled(param_dict) and get_eigenvalue_enabled(param_dict):
        return tuple(_eig(param_dict, k) for k, _ in _EIGEN)
    return tuple(v for _, v in _EIGEN)


def get_quantize_enabled(param_dict):
    """Quantize-aware training of the compression section (weight quantization with shared parameters enabled)."""
    wq = _sub(_sub(param_dict, "compression_training"), "weight_quantization")
    return bool(_p(_sub(wq, "shared_parameters"), "enabled", False))


def get_pld_enabled(param_dict):
    return _p(_sub(param_dict, "progressive_layer_drop"), "enabled", False)


def get_pld_params(param_dict):
    if "progressive_layer_drop" not in (param_dict or {}):
        return False
    params = copy.copy(param_dict["progressive_layer_drop"])
    params.pop("enabled", None)
    return params


# ---- data efficiency ---------------------------------------------------------------------------------------------------
def _de(param_dict):
    return _sub(param_dict, "data_efficiency")


def get_data_sampling_enabled(param_dict):
    return _p(_sub(_de(param_dict), "data_sampling"), "enabled", False)


def get_data_sampling_num_epochs(param_dict):
    return _p(_sub(_de(param_dict), "data_sampling"), "num_epochs", 1000)


def get_data_sampling_num_workers(param_dict):
    return _p(_sub(_de(param_dict), "data_sampling"), "num_workers", 0)


def get_data_sampling_pin_memory(param_dict):
    return _p(_sub(_de(param_dict), "data_sampling"), "pin_memory", False)


def get_curriculum_learning_enabled(param_dict):
    return _p(_sub(_sub(_de(param_dict), "data_sampling"), "curriculum_learning"), "enabled", False)


def get_curriculum_learning_params(param_dict):
    cl = copy.copy(_sub(_sub(_de(param_dict), "data_sampling"), "curriculum_learning"))
    cl.pop("enabled", None)
    return cl


def get_dynamic_batching(param_dict):
    d = {"enabled": False, "lr_scaling_method": "linear", "min_batch_size": 1, "max_batch_size": None,
         "sequence_picking_order": "dataloader", "verbose": False}
    d.update(_sub(_sub(_de(param_dict), "data_sampling"), "dynamic_batching"))
    return d


def get_data_routing_enabled(param_dict):
    return _p(_sub(_de(param_dict), "data_routing"), "enabled", False)


def get_random_ltd_enabled(param_dict):
    return _p(_sub(_sub(_de(param_dict), "data_routing"), "random_ltd"), "enabled", False)


def get_random_ltd_params(param_dict):
    ltd = copy.copy(_sub(_sub(_de(param_dict), "data_routing"), "random_ltd"))
    ltd.pop("enabled", None)
    return ltd
