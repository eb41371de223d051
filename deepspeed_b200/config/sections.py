"""All typed ds_config sections in one tree.

Key surface and defaults follow the reference: ZeRO ``runtime/zero/config.py:90-401``; offload
``runtime/zero/offload_config.py:20-88``; fp16/bf16 ``runtime/precision_config.py``; TP
``runtime/tensor_parallel/config.py:18-105``; comms logger ``comm/config.py``; monitor
``monitor/config.py``; flops profiler ``profiling/config.py``; activation checkpointing
``runtime/activation_checkpointing/config.py``; aio ``runtime/swap_tensor/aio_config.py``; checkpoint
``runtime/model_checkpointing/config.py``; compile ``compile/config.py``; elasticity
``elasticity/config.py``; data pipeline ``runtime/data_pipeline/config.py``.
"""
import sys
from enum import Enum
from pathlib import Path
from typing import Any, Dict, List, Optional, Union

from pydantic import Field, model_validator

from .base import DeepSpeedConfigModel, pp_int


# ------------------------------------------------------------------ ZeRO ----------------------
class ZeroStageEnum(int, Enum):
    disabled = 0
    optimizer_states = 1
    gradients = 2
    weights = 3
    max_stage = 3


class OffloadDeviceEnum(str, Enum):
    none = "none"
    cpu = "cpu"
    nvme = "nvme"


class OffloadStateTypeEnum(str, Enum):
    optim_states = "optim_states"
    hp_params = "hp_params"
    lp_params = "lp_params"
    lp_grads = "lp_grads"
    contiguous_grad_buffer = "contiguous_grad_buffer"


class DeepSpeedZeroOffloadParamConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = "none"
    nvme_path: Optional[Path] = None
    buffer_count: int = Field(5, ge=0)
    buffer_size: int = Field(pp_int(1e8), ge=0)
    max_in_cpu: int = Field(pp_int(1e9), ge=0)
    pin_memory: bool = True


class DeepSpeedZeroOffloadOptimizerConfig(DeepSpeedConfigModel):
    device: OffloadDeviceEnum = "none"
    nvme_path: Optional[Path] = None
    buffer_count: int = Field(4, ge=0)
    pin_memory: bool = True
    pipeline_read: bool = False
    pipeline_write: bool = False
    fast_init: bool = False
    ratio: float = Field(1.0, ge=0.0, le=1.0)
    super_offload: bool = False
    cpuadam_cores_perc: float = Field(0.8, ge=0.0, le=1.0)

    @property
    def pipeline(self):
        return self.pipeline_read or self.pipeline_write


class ZenFlowConfig(DeepSpeedConfigModel):
    topk_ratio: float = Field(0.1, ge=0.0, le=1.0)
    select_strategy: str = "auto"
    select_interval: Union[str, int] = "auto"
    update_interval: Union[str, int] = "auto"
    overlap_step: bool = False
    offload: bool = False
    auto_ratio: float = Field(0.99, ge=0.0, le=1.0)
    full_warm_up_rounds: int = 0
    steps_per_epoch: Optional[int] = None
    pt_reserved_cores_perc: float = Field(0.5, ge=0.0, le=1.0)


DEFAULT_LEAF_MODULE_CLASSES = [
    "transformers.models.mixtral.modeling_mixtral.MixtralSparseMoeBlock",
    "transformers.models.qwen2_moe.modeling_qwen2_moe.Qwen2MoeSparseMoeBlock",
    "transformers.models.qwen3_moe.modeling_qwen3_moe.Qwen3MoeSparseMoeBlock",
]


class DeepSpeedZeroLeafModuleConfig(DeepSpeedConfigModel):
    """``zero_optimization.leaf_module`` (reference ``runtime/zero/leaf_module_config.py``): HF sparse-MoE blocks are
    leaves by default; a bare string is accepted for each list; duplicates are dropped."""
    classes: List[str] = Field(default_factory=lambda: list(DEFAULT_LEAF_MODULE_CLASSES))
    names: List[str] = Field(default_factory=list)
    name_suffixes: List[str] = Field(default_factory=list)

    @model_validator(mode="before")
    @classmethod
    def _coerce(cls, values):
        if values is None:
            return {}
        if not isinstance(values, dict):
            raise TypeError("leaf_module configuration must be a mapping of fields to values")
        out = dict(values)
        for key in ("classes", "names", "name_suffixes"):
            if isinstance(out.get(key), str):
                out[key] = [out[key]]
        return out

    @model_validator(mode="after")
    def _dedupe(self):
        for key in ("classes", "names", "name_suffixes"):
            object.__setattr__(self, key, list(dict.fromkeys(str(v) for v in getattr(self, key))))
        return self


class DeepSpeedZeroConfig(DeepSpeedConfigModel):
    stage: ZeroStageEnum = 0
    contiguous_gradients: bool = True
    reduce_scatter: bool = True
    reduce_bucket_size: int = Field(pp_int(5e8), ge=0)
    use_multi_rank_bucket_allreduce: bool = True
    allgather_partitions: bool = True
    allgather_bucket_size: int = Field(pp_int(5e8), ge=0)
    overlap_comm: Optional[bool] = None  # None -> True for stage 3 else False
    load_from_fp32_weights: bool = True
    elastic_checkpoint: bool = False
    offload_param: Optional[DeepSpeedZeroOffloadParamConfig] = None
    offload_optimizer: Optional[DeepSpeedZeroOffloadOptimizerConfig] = None
    zenflow: Optional[ZenFlowConfig] = None
    sub_group_size: int = Field(pp_int(1e9), ge=0)
    cpu_offload_param: Optional[bool] = Field(
        None, json_schema_extra={
            "deprecated": True,
            "new_param": "offload_param",
            "new_param_fn": (lambda val: DeepSpeedZeroOffloadParamConfig(device="cpu") if val else None)
        })
    cpu_offload_use_pin_memory: Optional[bool] = Field(
        None, json_schema_extra={
            "deprecated": True,
            "new_param": "offload_param or offload_optimizer",
            "set_new_param": False
        })
    cpu_offload: Optional[bool] = Field(
        None, json_schema_extra={
            "deprecated": True,
            "new_param": "offload_optimizer",
            "new_param_fn": (lambda val: DeepSpeedZeroOffloadOptimizerConfig(device="cpu") if val else None)
        })
    prefetch_bucket_size: int = Field(pp_int(5e7), ge=0, alias="stage3_prefetch_bucket_size")
    param_persistence_threshold: int = Field(pp_int(1e5), ge=0, alias="stage3_param_persistence_threshold")
    model_persistence_threshold: int = Field(pp_int(sys.maxsize, "sys.maxsize"),
                                             ge=0,
                                             alias="stage3_model_persistence_threshold")
    max_live_parameters: int = Field(pp_int(1e9), ge=0, alias="stage3_max_live_parameters")
    max_reuse_distance: int = Field(pp_int(1e9), ge=0, alias="stage3_max_reuse_distance")
    gather_16bit_weights_on_model_save: bool = Field(False, alias="stage3_gather_16bit_weights_on_model_save")
    module_granularity_threshold: int = Field(pp_int(0), alias="stage3_module_granularity_threshold")
    use_all_reduce_for_fetch_params: bool = Field(False, alias="stage3_use_all_reduce_for_fetch_params")
    allgather_sequential: bool = Field(False, alias="stage3_allgather_sequential")
    stage3_gather_fp16_weights_on_model_save: bool = Field(False,
                                                           json_schema_extra={
                                                               "deprecated": True,
                                                               "new_param": "gather_16bit_weights_on_model_save"
                                                           })
    ignore_unused_parameters: bool = True
    legacy_stage1: bool = False
    round_robin_gradients: bool = False
    zero_hpz_partition_size: int = Field(1, ge=0)
    zero_quantized_weights: bool = False
    zero_quantized_nontrainable_weights: bool = False
    zero_quantized_gradients: bool = False
    zeropp_loco_param: Optional[Dict[str, Any]] = None
    mics_shard_size: int = Field(-1, json_schema_extra={"new_param": "mics_shard_size"})
    mics_hierarchical_params_gather: bool = False
    memory_efficient_linear: bool = True
    pipeline_loading_checkpoint: bool = False
    override_module_apply: bool = True
    log_trace_cache_warnings: bool = False
    enable_sanity_checks: bool = False
    save_muon_momentum_buffer_in_memory: bool = False
    leaf_module: DeepSpeedZeroLeafModuleConfig = Field(default_factory=DeepSpeedZeroLeafModuleConfig)
    # ---- B200-native knobs (no reference counterpart) ----
    fused_collectives: Optional[bool] = None  # None -> on when symmetric memory is available on NVLink
    fused_rs_adam: Optional[bool] = None  # None -> on when no clipping / loss scale / accumulation forces a late step
    unit_max_numel: int = Field(pp_int(1e9), ge=0)  # ZeRO-3 unit (flat shard) upper size in elements
    quantized_weights_format: str = "fp8_e4m3_block32"  # qwZ wire format: block-scaled fp8 or "int8"
    quantized_gradients_format: str = "fp8_e4m3_block32"

    @model_validator(mode="after")
    def overlap_comm_valid(self):
        if self.overlap_comm is None:
            object.__setattr__(self, "overlap_comm", self.stage == ZeroStageEnum.weights)
        return self

    @model_validator(mode="after")
    def offload_ratio_check(self):
        oc = self.offload_optimizer
        if oc is not None and oc.ratio < 1.0:
            assert self.stage == ZeroStageEnum.weights, "Partial offloading only supported for ZeRO Stage 3."
        return self


# ------------------------------------------------------------------ precision -----------------
class DeepSpeedFP16Config(DeepSpeedConfigModel):
    enabled: bool = False
    auto_cast: bool = False
    loss_scale: float = Field(0, ge=0)  # 0 -> dynamic
    initial_scale_power: int = Field(16, ge=0)
    loss_scale_window: int = Field(1000, ge=0)
    hysteresis: int = Field(2, ge=0)
    consecutive_hysteresis: bool = False
    min_loss_scale: float = Field(1, ge=0)
    fp16_master_weights_and_grads: bool = False
    fp16_opt_level: Optional[str] = None

    def initial_dynamic_scale(self):
        return 2**self.initial_scale_power

    def dynamic_loss_scale_args(self):
        return {
            "init_scale": 2**self.initial_scale_power,
            "scale_window": self.loss_scale_window,
            "min_scale": self.min_loss_scale,
            "delayed_shift": self.hysteresis,
            "consecutive_hysteresis": self.consecutive_hysteresis,
        }


class DeepSpeedBF16Config(DeepSpeedConfigModel):
    enabled: bool = False
    immediate_grad_update: bool = True
    check_grad_overflow: bool = False
    bf16_master_weights_and_grads: bool = False
    bf16_optimizer_states: bool = False


class DeepSpeedAMPConfig(DeepSpeedConfigModel):
    enabled: bool = False
    opt_level: Optional[str] = None
    model_config = dict(DeepSpeedConfigModel.model_config, extra="allow")


class TorchAutocastConfig(DeepSpeedConfigModel):
    enabled: bool = False
    dtype: str = "bfloat16"
    lower_precision_safe_modules: Optional[List[str]] = None


class DataTypesConfig(DeepSpeedConfigModel):
    grad_accum_dtype: Optional[str] = None
    param_dtype: Optional[str] = None
    buffer_dtype: Optional[str] = None


# ------------------------------------------------------------------ optimizer / scheduler ------
class OptimizerConfig(DeepSpeedConfigModel):
    type: Optional[str] = None
    params: Dict[str, Any] = Field(default_factory=dict)
    legacy_fusion: bool = False


class SchedulerConfig(DeepSpeedConfigModel):
    type: Optional[str] = None
    params: Dict[str, Any] = Field(default_factory=dict)


# ------------------------------------------------------------------ parallelism ----------------
class TPConfig(DeepSpeedConfigModel):
    tp_size: int = 1
    tp_grain_size: int = 1
    mpu: Any = None
    tp_group: Any = None


class TPTrainingConfig(DeepSpeedConfigModel):
    dtype: Any = "bfloat16"
    autotp_size: int = 0
    tp_overlap_comm: bool = False
    tensor_parallel: TPConfig = Field(default_factory=TPConfig, alias="tp")
    injection_policy_tuple: Optional[Any] = None
    keep_module_on_host: bool = False
    replace_with_kernel_inject: bool = False
    partition_config: Optional[Dict[str, Any]] = None
    preset_model: Optional[str] = None
    fused_collectives: bool = True  # B200: GEMM->reduce-scatter / all-gather->GEMM fused kernels


class ExpertParallelConfig(DeepSpeedConfigModel):
    """``expert_parallel`` section (reference ``module_inject/auto_ep_config.py::parse_autoep_config``); the keys the
    user actually wrote are what ``parallel/auto_ep_presets.parse_autoep_config`` hands to AutoEP."""
    enabled: bool = False
    autoep_size: int = 1
    expert_tensor_parallel_size: int = 1
    validate_folding_routing: bool = False
    preset_model: Optional[str] = None
    moe_layer_pattern: Optional[str] = None
    router_pattern: Optional[str] = None
    expert_pattern: Optional[str] = None
    num_experts: Optional[int] = None
    use_grouped_mm: bool = True
    disable_triton_grouped_mm: bool = False
    route_norm: Optional[bool] = None
    route_scale: float = 1.0
    score_apply: str = "auto"
    combine_impl: str = "auto"
    num_expert_groups: Optional[int] = None
    num_limited_groups: Optional[int] = None
    score_func: str = "auto"
    top_k: Any = "auto"
    load_balance_coeff: Optional[float] = None
    routed_scaling_factor: Any = "auto"
    expert_w1: Optional[str] = None
    expert_w2: Optional[str] = None
    expert_w3: Optional[str] = None
    num_experts_attr: Optional[str] = None
    top_k_attr: Optional[str] = None
    has_shared_experts: Optional[bool] = None
    shared_experts_pattern: Optional[str] = None
    shared_experts_gate_pattern: Optional[str] = None
    return_router_logits: Optional[bool] = None
    fused_dispatch: bool = True  # B200: in-kernel NVLink dispatch/combine
    model_config = dict(DeepSpeedConfigModel.model_config, extra="allow")


class PipelineConfig(DeepSpeedConfigModel):
    stages: Union[str, int] = "auto"
    partition: str = "best"
    seed_layers: bool = False
    activation_checkpoint_interval: int = 0
    pipe_partitioned: bool = True
    grad_partitioned: bool = True
    use_reentrant: bool = True


# ------------------------------------------------------------------ aux ------------------------
class CommsLoggerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    prof_all: bool = True
    prof_ops: List[str] = Field(default_factory=list)
    verbose: bool = False
    debug: bool = False


class TensorBoardConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class WandbConfig(DeepSpeedConfigModel):
    enabled: bool = False
    group: Optional[str] = None
    team: Optional[str] = None
    project: str = "deepspeed"


class CSVConfig(DeepSpeedConfigModel):
    enabled: bool = False
    output_path: str = ""
    job_name: str = "DeepSpeedJobName"


class CometConfig(DeepSpeedConfigModel):
    enabled: bool = False
    samples_log_interval: int = 100
    project: Optional[str] = None
    workspace: Optional[str] = None
    api_key: Optional[str] = None
    experiment_name: Optional[str] = None
    experiment_key: Optional[str] = None
    online: Optional[bool] = None
    mode: Optional[str] = None


class DeepSpeedMonitorConfig(DeepSpeedConfigModel):
    tensorboard: TensorBoardConfig = Field(default_factory=TensorBoardConfig)
    comet: CometConfig = Field(default_factory=CometConfig)
    wandb: WandbConfig = Field(default_factory=WandbConfig)
    csv_monitor: CSVConfig = Field(default_factory=CSVConfig)

    @property
    def enabled(self):
        return self.tensorboard.enabled or self.wandb.enabled or self.csv_monitor.enabled or self.comet.enabled


class DeepSpeedFlopsProfilerConfig(DeepSpeedConfigModel):
    enabled: bool = False
    recompute_fwd_factor: float = 0.0
    profile_step: int = 1
    module_depth: int = -1
    top_modules: int = 1
    detailed: bool = True
    output_file: Optional[str] = None


class DeepSpeedActivationCheckpointingConfig(DeepSpeedConfigModel):
    partition_activations: bool = False
    contiguous_memory_optimization: bool = False
    cpu_checkpointing: bool = False
    number_checkpoints: Optional[int] = None
    synchronize_checkpoint_boundary: bool = False
    profile: bool = False


class DeepSpeedAIOConfig(DeepSpeedConfigModel):
    block_size: int = Field(pp_int(1048576), ge=0)
    queue_depth: int = Field(8, ge=0)
    intra_op_parallelism: int = Field(1, ge=0, alias="thread_count")
    single_submit: bool = False
    overlap_events: bool = True
    use_gds: bool = False


class CheckpointWriterConfig(DeepSpeedConfigModel):
    type: str = "torch"  # torch | fast | decoupled
    io_buffer_size: int = Field(pp_int(1024 * 1024 * 1024), ge=0)
    io_buffer_double: bool = True
    io_multiplier: int = 1
    io_statistics: bool = False
    show_statistics: bool = False
    data_parallel: Optional[str] = None
    decoupled: bool = False
    model_config = dict(DeepSpeedConfigModel.model_config, extra="allow")


class CheckpointConfig(DeepSpeedConfigModel):
    tag_validation: str = "Warn"  # Ignore | Warn | Fail
    load_universal: bool = False
    use_node_local_storage: bool = False
    parallel_write: Dict[str, Any] = Field(default_factory=dict)
    writer: Optional[CheckpointWriterConfig] = None

    @model_validator(mode="after")
    def _check_tag(self):
        mode = str(self.tag_validation).capitalize()
        assert mode in ("Ignore", "Warn", "Fail"), f"checkpoint.tag_validation must be Ignore|Warn|Fail, got {mode}"
        object.__setattr__(self, "tag_validation", mode)
        return self


class CompileConfig(DeepSpeedConfigModel):
    deepcompile: bool = False
    free_activation: bool = False
    free_activation_threshold: int = 10 * 1024 * 1024
    offload_activation: bool = False
    offload_activation_pin_memory: bool = True
    offload_opt_states: bool = False
    double_buffer: bool = True
    symmetric_memory: bool = False
    debug_log: bool = False
    offload_parameters: bool = False
    sync_before_reduce: bool = False
    sync_after_reduce: bool = False
    sync_before_allgather: bool = False
    sync_after_allgather: bool = False
    keep_int_input_tensors: bool = True
    keep_all_input_tensors: bool = False
    passes: Optional[List[str]] = None
    model_config = dict(DeepSpeedConfigModel.model_config, extra="allow")


class ElasticityConfigModel(DeepSpeedConfigModel):
    enabled: bool = False
    max_train_batch_size: int = 2000
    micro_batch_sizes: List[int] = Field(default_factory=lambda: [2, 4, 6])
    min_gpus: int = 1
    max_gpus: int = 10000
    min_time: int = 0
    version: float = 0.2
    prefer_larger_batch: bool = True
    ignore_non_elastic_batch_info: bool = False
    num_gpus_per_node: int = 1
    model_parallel_size: int = 1


class HybridEngineConfig(DeepSpeedConfigModel):
    enabled: bool = False
    max_out_tokens: int = 512
    inference_tp_size: int = 1
    release_inference_cache: bool = False
    pin_parameters: bool = True
    tp_gather_partition_size: int = 8
    enable_cuda_graph: bool = False  # replay the decode forward from CUDA graphs (runtime/hybrid_engine_graph.py)


class GraphHarvestingConfig(DeepSpeedConfigModel):
    enabled: bool = False


class TimersThroughputConfig(DeepSpeedConfigModel):
    enabled: bool = True
    synchronized: bool = True


class TimersConfig(DeepSpeedConfigModel):
    throughput: TimersThroughputConfig = Field(default_factory=TimersThroughputConfig)


# ---- reference helpers of ``runtime/zero/config.py`` ---------------------------------------------------------
ZERO_OPTIMIZATION = "zero_optimization"


def read_zero_config_deprecated(param_dict):
    """``"zero_optimization": true`` (pre-0.3 format): stage 1, bucket size from ``allgather_size``."""
    import logging
    cfg = {"stage": 1 if param_dict[ZERO_OPTIMIZATION] else 0}
    if cfg["stage"] > 0:
        cfg["allgather_bucket_size"] = param_dict.get("allgather_size", 5e8)
    logging.getLogger("DeepSpeed").warning(
        "DeepSpeedConfig: this format of ZeRO optimization setup is deprecated. Please use "
        '"zero_optimization": {"stage": [0|1|2|3], ...}')
    return cfg


def get_zero_config(param_dict):
    section = param_dict.get(ZERO_OPTIMIZATION, {})
    if isinstance(section, bool):
        section = read_zero_config_deprecated(param_dict)
    return DeepSpeedZeroConfig(**section)
