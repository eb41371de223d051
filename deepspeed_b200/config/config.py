"""Top-level ``DeepSpeedConfig``: JSON/hjson/dict/base64 -> typed tree, batch-size triad, sanity checks.

Reference behaviour: ``deepspeed/runtime/config.py:692-1060`` (``DeepSpeedConfig``,
``_initialize_params`` :796-923, ``_set_batch_related_parameters`` :942-987, ``_do_error_check`` :1017),
key constants ``deepspeed/runtime/constants.py``. The section models are in :mod:`.sections`.
"""
import base64
import copy
import json
import os
from typing import Optional

from ..utils.logging import logger, set_log_level
from .base import (DeepSpeedConfigObject, ScientificNotationEncoder, dict_raise_error_on_duplicate_keys,
                   get_scalar_param)
from .sections import (CheckpointConfig, CommsLoggerConfig, CompileConfig, DataTypesConfig,
                       DeepSpeedActivationCheckpointingConfig, DeepSpeedAIOConfig, DeepSpeedAMPConfig,
                       DeepSpeedBF16Config, DeepSpeedFlopsProfilerConfig, DeepSpeedFP16Config, DeepSpeedMonitorConfig,
                       DeepSpeedZeroConfig, ElasticityConfigModel, ExpertParallelConfig, GraphHarvestingConfig,
                       HybridEngineConfig, OptimizerConfig, PipelineConfig, SchedulerConfig, TimersConfig,
                       TorchAutocastConfig, TPTrainingConfig, ZeroStageEnum)

# ---- key constants (same spellings as the reference's runtime/constants.py) ----
TRAIN_BATCH_SIZE = "train_batch_size"
TRAIN_MICRO_BATCH_SIZE_PER_GPU = "train_micro_batch_size_per_gpu"
GRADIENT_ACCUMULATION_STEPS = "gradient_accumulation_steps"
ZERO_OPTIMIZATION = "zero_optimization"
GRADIENT_CLIPPING = "gradient_clipping"
STEPS_PER_PRINT = "steps_per_print"

ADAGRAD_OPTIMIZER = "adagrad"
ADAM_OPTIMIZER = "adam"
ADAMW_OPTIMIZER = "adamw"
LAMB_OPTIMIZER = "lamb"
ONEBIT_ADAM_OPTIMIZER = "onebitadam"
ZERO_ONE_ADAM_OPTIMIZER = "zerooneadam"
ONEBIT_LAMB_OPTIMIZER = "onebitlamb"
MUADAM_OPTIMIZER = "muadam"
MUADAMW_OPTIMIZER = "muadamw"
MUSGD_OPTIMIZER = "musgd"
LION_OPTIMIZER = "lion"
MUON_OPTIMIZER = "muon"
SGD_OPTIMIZER = "sgd"
DEEPSPEED_OPTIMIZERS = [
    ADAGRAD_OPTIMIZER, ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER, ONEBIT_LAMB_OPTIMIZER,
    ZERO_ONE_ADAM_OPTIMIZER, MUADAM_OPTIMIZER, MUADAMW_OPTIMIZER, MUSGD_OPTIMIZER, LION_OPTIMIZER, MUON_OPTIMIZER,
    SGD_OPTIMIZER
]
TORCH_ADAM_PARAM = "torch_adam"
ADAM_W_MODE = "adam_w_mode"
ADAM_W_MODE_DEFAULT = True

_SECTION_KEYS_HANDLED = {
    TRAIN_BATCH_SIZE, TRAIN_MICRO_BATCH_SIZE_PER_GPU, GRADIENT_ACCUMULATION_STEPS, "managed_gradient_accumulation",
    STEPS_PER_PRINT, "optimizer", "scheduler", "fp16", "bf16", "bfloat16", "amp", "torch_autocast", GRADIENT_CLIPPING,
    "prescale_gradients", "gradient_predivide_factor", "gradient_allreduce_op", "communication_data_type",
    "seq_parallel_communication_data_type", "sparse_gradients", ZERO_OPTIMIZATION, "zero_allow_untested_optimizer",
    "zero_force_ds_cpu_optimizer", "activation_checkpointing", "aio", "pipeline", "tensor_parallel",
    "expert_parallel", "sequence_parallel_size", "data_parallel_size", "checkpoint", "data_types", "comms_logger",
    "tensorboard", "wandb", "csv_monitor", "comet", "flops_profiler", "wall_clock_breakdown", "memory_breakdown",
    "elasticity", "autotuning", "compression_training", "data_efficiency", "curriculum_learning",
    "progressive_layer_drop", "eigenvalue", "sparse_attention", "weight_quantization", "hybrid_engine", "compile",
    "nebula", "datastates", "timers", "graph_harvesting", "dataloader_drop_last", "use_data_before_expert_parallelism",
    "vocabulary_size", "dump_state", "log_level", "disable_allgather", "zero_enabled", "use_node_local_storage",
    "load_universal_checkpoint", "quantize_training", "train_micro_batch_size", "torch_adam", "float16",
}


class DeepSpeedConfigError(Exception):
    pass


def load_ds_config(config):
    """Accept a path (json/hjson), a dict, or a base64-encoded JSON string; return a dict."""
    if isinstance(config, dict):
        return copy.deepcopy(config)
    if isinstance(config, (str, os.PathLike)):
        config = str(config)
        if os.path.exists(config):
            with open(config) as f:
                text = f.read()
            try:
                return json.loads(text, object_pairs_hook=dict_raise_error_on_duplicate_keys)
            except json.JSONDecodeError:
                import hjson
                return hjson.loads(text, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        try:
            decoded = base64.urlsafe_b64decode(config).decode("utf-8")
            return json.loads(decoded)
        except Exception:
            raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary or a valid "
                             f"base64. Received: {config}")
    raise ValueError(f"Unsupported ds_config type {type(config)}")


def _dtype_from_str(name):
    import torch
    if name is None:
        return None
    table = {
        "fp32": torch.float32, "float32": torch.float32, "float": torch.float32, "fp16": torch.float16,
        "float16": torch.float16, "half": torch.float16, "bf16": torch.bfloat16, "bfloat16": torch.bfloat16,
        "int8": torch.int8, "fp8": getattr(torch, "float8_e4m3fn", None),
    }
    if name not in table:
        raise ValueError(f"Invalid data type '{name}'. Supported: {sorted(table)}")
    return table[name]


class DeepSpeedConfig:
    """Parsed ds_config. Attribute names follow the reference so downstream code reads the same."""

    def __init__(self, config, mpu=None, mesh_device=None, world_size: Optional[int] = None):
        self._param_dict = load_ds_config(config)
        pd = self._param_dict
        # ---- world size for the batch triad: data-parallel replicas only ----
        from .. import comm as dist
        if world_size is not None:
            self.world_size = world_size
            self.global_rank = dist.get_rank() if dist.is_initialized() else 0
        else:
            try:
                self.global_rank = dist.get_rank()
                total = dist.get_world_size()
            except Exception:
                self.global_rank, total = 0, 1
            if mpu is not None and hasattr(mpu, "get_data_parallel_world_size"):
                self.world_size = mpu.get_data_parallel_world_size()
            elif mesh_device is not None:
                self.world_size = mesh_device.get_group("data_parallel").size() if hasattr(
                    mesh_device, "get_group") else dist.get_world_size(mesh_device.get_group(mesh_dim="data_parallel"))
            else:
                sp = int(pd.get("sequence_parallel_size", 1) or 1)
                tp_cfg = pd.get("tensor_parallel", {}) or {}
                tp = int(tp_cfg.get("autotp_size", 0) or 0) or int((tp_cfg.get("tp", {}) or {}).get("tp_size", 1) or 1)
                pp = pd.get("pipeline", {}).get("stages", 1) if isinstance(pd.get("pipeline"), dict) else 1
                pp = pp if isinstance(pp, int) else 1
                self.world_size = max(1, total // max(1, sp * tp * pp))
        # ---- elasticity may override the batch triad (reference config.py:730-790) ----
        self.elasticity_enabled = bool(pd.get("elasticity", {}).get("enabled", False))
        if self.elasticity_enabled:
            from ..elasticity import apply_elasticity_to_config
            apply_elasticity_to_config(pd, self.world_size)
        self._initialize_params(pd)
        self._configure_train_batch_size()
        self._do_sanity_check()

    # ------------------------------------------------------------------------------------------
    def _initialize_params(self, pd):
        unknown = set(pd.keys()) - _SECTION_KEYS_HANDLED
        if unknown:
            logger.warning(f"ds_config: ignoring unknown top-level keys {sorted(unknown)}")
        self.train_batch_size = pd.get(TRAIN_BATCH_SIZE, None)
        self.train_micro_batch_size_per_gpu = pd.get(TRAIN_MICRO_BATCH_SIZE_PER_GPU, None)
        self.gradient_accumulation_steps = pd.get(GRADIENT_ACCUMULATION_STEPS, None)
        self.managed_gradient_accumulation = pd.get("managed_gradient_accumulation", True)
        self.steps_per_print = pd.get(STEPS_PER_PRINT, None)
        self.dump_state = pd.get("dump_state", False)
        self.disable_allgather = pd.get("disable_allgather", False)
        self.communication_data_type = _dtype_from_str(pd.get("communication_data_type", None))
        self.seq_parallel_communication_data_type = _dtype_from_str(
            pd.get("seq_parallel_communication_data_type", "fp32"))
        self.prescale_gradients = pd.get("prescale_gradients", False)
        self.gradient_predivide_factor = pd.get("gradient_predivide_factor", 1.0)
        self.gradient_allreduce_op = str(pd.get("gradient_allreduce_op", "mean")).lower()
        assert self.gradient_allreduce_op in ("mean", "sum"), "gradient_allreduce_op must be 'mean' or 'sum'"
        self.sparse_gradients_enabled = pd.get("sparse_gradients", False)

        zero_dict = pd.get(ZERO_OPTIMIZATION, {})
        if isinstance(zero_dict, bool):  # legacy "zero_optimization": true -> stage 1
            zero_dict = {"stage": 1 if zero_dict else 0}
        self.zero_config = DeepSpeedZeroConfig(**zero_dict)
        self.zero_optimization_stage = int(self.zero_config.stage)
        self.zero_enabled = self.zero_optimization_stage > 0
        self.zero_allow_untested_optimizer = pd.get("zero_allow_untested_optimizer", False)
        self.zero_force_ds_cpu_optimizer = pd.get("zero_force_ds_cpu_optimizer", True)
        self.mics_shard_size = self.zero_config.mics_shard_size
        self.mics_hierarchial_params_gather = self.zero_config.mics_hierarchical_params_gather

        self.activation_checkpointing_config = DeepSpeedActivationCheckpointingConfig(
            **pd.get("activation_checkpointing", {}))
        self.comms_config = self.comms_logger = CommsLoggerConfig(**pd.get("comms_logger", {}))
        self.monitor_config = DeepSpeedMonitorConfig(
            **{k: pd[k] for k in ("tensorboard", "wandb", "csv_monitor", "comet") if k in pd})
        self.flops_profiler_config = DeepSpeedFlopsProfilerConfig(**pd.get("flops_profiler", {}))
        self.aio_config = DeepSpeedAIOConfig(**pd.get("aio", {}))
        self.compile_config = CompileConfig(**pd.get("compile", {}))
        self.timers_config = TimersConfig(**pd.get("timers", {}))
        self.tensor_parallel_config = TPTrainingConfig(**(pd.get("tensor_parallel", {}) or {}))
        self.expert_parallel_config = ExpertParallelConfig(**(pd.get("expert_parallel", {}) or {}))
        self.hybrid_engine = HybridEngineConfig(**pd.get("hybrid_engine", {}))
        self.graph_harvesting = bool(GraphHarvestingConfig(**pd["graph_harvesting"]).enabled) if isinstance(
            pd.get("graph_harvesting"), dict) else bool(pd.get("graph_harvesting", False))
        self.sequence_parallel_size = int(pd.get("sequence_parallel_size", 1) or 1)
        self.data_parallel_size = pd.get("data_parallel_size", None)

        self.gradient_clipping = float(pd.get(GRADIENT_CLIPPING, 0.0))
        fp16_dict = pd.get("fp16", pd.get("float16", {}))
        self.float16_config = DeepSpeedFP16Config(**fp16_dict)
        self.bfloat16_config = DeepSpeedBF16Config(**pd.get("bf16", pd.get("bfloat16", {})))
        assert not (self.float16_config.enabled
                    and self.bfloat16_config.enabled), "bfloat16 and fp16 modes cannot be simultaneously enabled"
        self.fp16_enabled = self.float16_config.enabled
        self.bfloat16_enabled = self.bfloat16_config.enabled
        self.fp16_auto_cast = self.float16_config.auto_cast
        self.fp16_master_weights_and_gradients = self.float16_config.fp16_master_weights_and_grads
        self.loss_scale = self.float16_config.loss_scale
        self.initial_dynamic_scale = self.float16_config.initial_dynamic_scale()
        self.dynamic_loss_scale_args = self.float16_config.dynamic_loss_scale_args()
        self.amp_config = DeepSpeedAMPConfig(**pd.get("amp", {}))
        self.amp_enabled = self.amp_config.enabled
        self.amp_params = {k: v for k, v in pd.get("amp", {}).items() if k != "enabled"}
        self.torch_autocast_config = TorchAutocastConfig(**pd.get("torch_autocast", {}))
        dt = DataTypesConfig(**pd.get("data_types", {}))
        self.data_types_config = dt
        self.grad_accum_dtype = dt.grad_accum_dtype
        self.param_dtype = dt.param_dtype

        opt = OptimizerConfig(**pd["optimizer"]) if isinstance(pd.get("optimizer"), dict) else OptimizerConfig()
        self.optimizer_name = opt.type.lower() if (opt.type and opt.type.lower() in DEEPSPEED_OPTIMIZERS) else opt.type
        self.optimizer_params = dict(opt.params)
        self.optimizer_legacy_fusion = opt.legacy_fusion
        sch = SchedulerConfig(**pd["scheduler"]) if isinstance(pd.get("scheduler"), dict) else SchedulerConfig()
        self.scheduler_name = sch.type
        self.scheduler_params = dict(sch.params)

        self.wall_clock_breakdown = bool(pd.get("wall_clock_breakdown", False)) | self.flops_profiler_config.enabled
        self.memory_breakdown = pd.get("memory_breakdown", False)
        self.eigenvalue_config = dict(pd.get("eigenvalue", {}))
        self.eigenvalue_enabled = bool(self.eigenvalue_config.get("enabled", False))
        self.pld_config = dict(pd.get("progressive_layer_drop", {}))
        self.pld_enabled = bool(self.pld_config.get("enabled", False))
        self.pld_params = {k: v for k, v in self.pld_config.items() if k != "enabled"} if self.pld_enabled else False
        self.curriculum_config = dict(pd.get("curriculum_learning", {}))
        self.curriculum_enabled_legacy = bool(self.curriculum_config.get("enabled", False))
        self.curriculum_params_legacy = self.curriculum_config if self.curriculum_enabled_legacy else False
        self.data_efficiency_config = dict(pd.get("data_efficiency", {}))
        self.data_efficiency_enabled = bool(self.data_efficiency_config.get("enabled", False))
        self.compression_config = dict(pd.get("compression_training", {}))
        self.sparse_attention = pd.get("sparse_attention", None)
        self.weight_quantization_config = pd.get("weight_quantization", None)
        self.quantize_training_config = pd.get("quantize_training", None)
        self.autotuning_config = dict(pd.get("autotuning", {}))
        self.elasticity_config = ElasticityConfigModel(**pd["elasticity"]) if "elasticity" in pd else None
        self.pipeline = PipelineConfig(**pd["pipeline"]).model_dump() if isinstance(pd.get("pipeline"), dict) \
            else PipelineConfig().model_dump()
        ck = dict(pd.get("checkpoint", {}))
        if "load_universal_checkpoint" in pd:
            ck.setdefault("load_universal", pd["load_universal_checkpoint"])
        if "use_node_local_storage" in pd:
            ck.setdefault("use_node_local_storage", pd["use_node_local_storage"])
        self.checkpoint_config = CheckpointConfig(**ck)
        self.checkpoint_tag_validation_enabled = self.checkpoint_config.tag_validation != "Ignore"
        self.checkpoint_tag_validation_fail = self.checkpoint_config.tag_validation == "Fail"
        self.load_universal_checkpoint = self.checkpoint_config.load_universal
        self.use_node_local_storage = self.checkpoint_config.use_node_local_storage
        self.checkpoint_parallel_write_pipeline = bool(
            self.checkpoint_config.parallel_write.get("pipeline_stage", False))
        self.nebula_config = dict(pd.get("nebula", {}))
        self.datastates_config = dict(pd.get("datastates", {}))
        self.dataloader_drop_last = pd.get("dataloader_drop_last", False)
        self.use_data_before_expert_parallel_ = pd.get("use_data_before_expert_parallelism", False)
        self.vocabulary_size = pd.get("vocabulary_size", None)
        if "log_level" in pd:
            set_log_level(pd["log_level"])

    # ------------------------------------------------------------------------------------------
    def _batch_assertion(self):
        train_batch, micro_batch, grad_acc = (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                                              self.gradient_accumulation_steps)
        assert train_batch > 0, f"Train batch size: {train_batch} has to be greater than 0"
        assert micro_batch > 0, f"Micro batch size per gpu: {micro_batch} has to be greater than 0"
        assert grad_acc > 0, f"Gradient accumulation steps: {grad_acc} has to be greater than 0"
        assert train_batch == micro_batch * grad_acc * self.world_size, (
            f"Check batch related parameters. train_batch_size is not equal to micro_batch_per_gpu * "
            f"gradient_acc_step * world_size {train_batch} != {micro_batch} * {grad_acc} * {self.world_size}")

    def _set_batch_related_parameters(self):
        """Any two of the triad determine the third (invariant: train = micro * gas * dp_world)."""
        train_batch, micro_batch, grad_acc = (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                                              self.gradient_accumulation_steps)
        ws = self.world_size
        if train_batch is not None and micro_batch is not None and grad_acc is not None:
            return
        if train_batch is not None and micro_batch is not None:
            grad_acc = train_batch // micro_batch
            grad_acc //= ws
            self.gradient_accumulation_steps = grad_acc
        elif train_batch is not None and grad_acc is not None:
            micro_batch = train_batch // ws
            micro_batch //= grad_acc
            self.train_micro_batch_size_per_gpu = micro_batch
        elif micro_batch is not None and grad_acc is not None:
            self.train_batch_size = micro_batch * grad_acc * ws
        elif train_batch is not None:
            self.gradient_accumulation_steps = 1
            self.train_micro_batch_size_per_gpu = train_batch // ws
        elif micro_batch is not None:
            self.train_batch_size = micro_batch * ws
            self.gradient_accumulation_steps = 1
        else:
            raise DeepSpeedConfigError("Either train_batch_size or train_micro_batch_size_per_gpu needs to be provided")

    def _configure_train_batch_size(self):
        self._set_batch_related_parameters()
        self._batch_assertion()

    def _do_sanity_check(self):
        self._do_error_check()
        self._do_warning_check()

    def _do_error_check(self):
        assert self.train_micro_batch_size_per_gpu, f"DeepSpeedConfig: {TRAIN_MICRO_BATCH_SIZE_PER_GPU} is not defined"
        assert self.gradient_accumulation_steps, f"DeepSpeedConfig: {GRADIENT_ACCUMULATION_STEPS} is not defined"
        if self.zero_enabled:
            assert self.zero_optimization_stage <= ZeroStageEnum.max_stage.value, \
                f"DeepSpeedConfig: Maximum supported ZeRO stage is {ZeroStageEnum.max_stage.value}"
        if self.fp16_master_weights_and_gradients:
            assert self.zero_enabled and self.zero_optimization_stage in (1, 2, 3), \
                "fp16_master_weights_and_grads is only supported with ZeRO"

    def _do_warning_check(self):
        fp16_enabled = self.fp16_enabled
        vocab = self.vocabulary_size
        if vocab and vocab % 8 != 0:
            logger.warning(f"DeepSpeedConfig: vocabulary size {vocab} is not aligned to 8, may import tensor core "
                           f"utilization.")
        if self.optimizer_params and "max_grad_norm" in self.optimizer_params and fp16_enabled:
            logger.warning("DeepSpeedConfig: In FP16 mode, DeepSpeed will pass max_grad_norm:"
                           f"{self.optimizer_params['max_grad_norm']} to FP16 wrapper")

    # ------------------------------------------------------------------------------------------
    def print_user_config(self):
        logger.info("  json = {}".format(
            json.dumps(self._param_dict, sort_keys=True, indent=4, cls=ScientificNotationEncoder, default=str)))

    def print(self, name):
        logger.info(f"{name}:")
        for arg in sorted(vars(self)):
            if arg != "_param_dict":
                dots = "." * max(1, 29 - len(arg))
                logger.info(f"  {arg} {dots} {getattr(self, arg)}")
        self.print_user_config()

    def to_dict(self):
        return copy.deepcopy(self._param_dict)


class DeepSpeedConfigWriter:
    """Programmatic config builder (reference ``runtime/config.py`` DeepSpeedConfigWriter)."""

    def __init__(self, data=None):
        self.data = data if data is not None else {}

    def add_config(self, key, value):
        self.data[key] = value

    def load_config(self, filename):
        self.data = load_ds_config(filename)

    def write_config(self, filename):
        with open(filename, "w") as outfile:
            json.dump(self.data, outfile)


__all__ = ["DeepSpeedConfig", "DeepSpeedConfigError", "DeepSpeedConfigWriter", "load_ds_config", "get_scalar_param",
           "DeepSpeedConfigObject"]
from .getters import *  # noqa: F401,F403,E402  (dictionary readers under the reference's names)
