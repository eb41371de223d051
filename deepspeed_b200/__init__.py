"""deepspeed_b200 -- a B200-native (sm_100a, NVLink 5 / NVSwitch) ZeRO training framework.

Public API mirrors the reference ``deepspeed/__init__.py``: ``initialize`` (:93-267), ``init_inference`` (:328),
``tp_model_init`` (:408), ``add_config_arguments`` (:305), ``init_distributed`` (re-export :56), ``zero.Init`` /
``zero.GatheredParameters``, ``PipelineModule``, ``checkpointing``, ``moe.layer.MoE``.
"""
import argparse
import os
import sys
import types
from typing import Optional, Union

import torch

from . import comm
from . import comm as dist
from .comm import init_distributed
from .config import DeepSpeedConfig, DeepSpeedConfigError, load_ds_config
from .runtime import zero
from .runtime.engine import DeepSpeedEngine, DeepSpeedOptimizerCallable, DeepSpeedSchedulerCallable
from .runtime import lr_schedules
from .runtime.lr_schedules import add_tuning_arguments
from .utils import groups, logging
from .utils.logging import log_dist, logger
from .version import __version__, __reference_api__ as _reference_api

__version_major__, __version_minor__, __version_patch__ = (int(x) for x in __version__.split(".")[:3])
__git_hash__ = None
__git_branch__ = None
git_hash, git_branch = __git_hash__, __git_branch__
version = __version__
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500
_TP_MODEL_INIT_ARGS = None


def _lazy(name):
    import importlib
    return importlib.import_module(f"{__name__}.{name}")


def __getattr__(name):
    lazy = {
        "moe": "moe", "ops": "ops", "pipe": "runtime.pipe", "checkpointing": "runtime.activation_checkpointing",
        "module_inject": "parallel", "sequence": "sequence", "monitor": "monitor", "profiling": "profiling",
        "elasticity": "elasticity", "launcher": "launcher", "checkpoint": "checkpoint", "compression": "compression",
        "inference": "inference", "linear": "linear", "autotuning": "autotuning", "models": "models",
        "parallel": "parallel", "accelerator": "accelerator", "io": "io", "nvme": "nvme",
    }
    if name in lazy:
        return _lazy(lazy[name])
    if name == "PipelineModule":
        return _lazy("runtime.pipe").PipelineModule
    if name == "PipelineEngine":
        return _lazy("runtime.pipe.engine").PipelineEngine
    if name == "InferenceEngine":
        return _lazy("inference.engine").InferenceEngine
    if name == "DeepSpeedTransformerLayer":
        return _lazy("ops.transformer").DeepSpeedTransformerLayer
    if name == "DeepSpeedTransformerConfig":
        return _lazy("ops.transformer").DeepSpeedTransformerConfig
    if name == "DeepSpeedHybridEngine":
        return _lazy("runtime.hybrid_engine").DeepSpeedHybridEngine
    if name == "get_accelerator":
        return _lazy("accelerator").get_accelerator
    late = {
        "OnDevice": ("utils.init_on_device", "OnDevice"),
        "DeepSpeedDataLoader": ("runtime.dataloader", "DeepSpeedDataLoader"),
        "DeepSpeedInferenceConfig": ("inference.config", "DeepSpeedInferenceConfig"),
        "DeepSpeedOptimizer": ("runtime.base_optimizer", "DeepSpeedOptimizer"),
        "replace_transformer_layer": ("parallel.replace_module", "replace_transformer_layer"),
        "revert_transformer_layer": ("parallel.replace_module", "revert_transformer_layer"),
        "set_autotp_mode": ("parallel.layers", "set_autotp_mode"),
        "is_compile_supported": ("runtime.compiler", "is_compile_supported"),
        "ADAM_OPTIMIZER": ("config.config", "ADAM_OPTIMIZER"),
        "LAMB_OPTIMIZER": ("config.config", "LAMB_OPTIMIZER"),
        "MUON_OPTIMIZER": ("config.config", "MUON_OPTIMIZER"),
    }
    if name in late:
        mod, attr = late[name]
        return getattr(_lazy(mod), attr)
    if name == "domino":
        return _lazy("runtime.domino")
    if name == "constants":
        return _lazy("constants")
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def initialize(args=None,
               model: torch.nn.Module = None,
               optimizer=None,
               model_parameters=None,
               training_data=None,
               lr_scheduler=None,
               distributed_port: int = 29500,
               mpu=None,
               dist_init_required: Optional[bool] = None,
               collate_fn=None,
               config=None,
               mesh_param=None,
               config_params=None):
    """Build the engine. Returns ``(engine, optimizer, training_dataloader, lr_scheduler)``."""
    log_dist(f"DeepSpeed-B200 info: version={__version__} (reference API {_reference_api})",
             ranks=[0])
    assert model is not None, "deepspeed.initialize requires a model"
    zero.shutdown_init_context()
    try:
        dist.init_distributed(distributed_port=distributed_port, dist_init_required=dist_init_required)
        if config is None:
            config = config_params
        if config is None and args is not None:
            config = getattr(args, "deepspeed_config", None) or getattr(args, "deepscale_config", None)
        assert config is not None, "DeepSpeed requires --deepspeed_config to specify configuration file"
        if hasattr(args, "deepspeed_config") and args.deepspeed_config is not None and config is not args.deepspeed_config \
                and config_params is None and not isinstance(config, (dict, str, os.PathLike)):
            raise ValueError("Not sure how to proceed, we were given deepspeed configs in the deepspeed arguments and "
                             "deepspeed.initialize() function call")
        cfg_dict = load_ds_config(config)
        mesh_device = None
        if mesh_param:
            log_dist(f"mesh_param to Initialize mesh device: {mesh_param}", [0])
            mesh_device = dist.init_device_mesh(tuple(mesh_param), ("data_parallel", "sequence_parallel"))
            cfg_dict.setdefault("sequence_parallel_size", mesh_param[1])
        elif "sequence_parallel_size" in cfg_dict and "data_parallel_size" in cfg_dict:
            cfg_dict.setdefault("sequence_parallel_size", cfg_dict["sequence_parallel_size"])
        merge_tp_model_init_into_config(cfg_dict, mpu, mesh_param, dist)
        config_class = DeepSpeedConfig(cfg_dict, mpu, mesh_device)
        from .runtime.pipe.module import PipelineModule
        if isinstance(model, PipelineModule):
            from .runtime.pipe.engine import PipelineEngine
            cfg_pp = DeepSpeedConfig(cfg_dict, model.mpu(), mesh_device)
            engine = PipelineEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                    training_data=training_data, lr_scheduler=lr_scheduler, mpu=model.mpu(),
                                    dist_init_required=dist_init_required, collate_fn=collate_fn, config=cfg_dict,
                                    config_class=cfg_pp)
        elif config_class.hybrid_engine.enabled:
            from .runtime.hybrid_engine import DeepSpeedHybridEngine
            engine = DeepSpeedHybridEngine(args=args, model=model, optimizer=optimizer,
                                           model_parameters=model_parameters, training_data=training_data,
                                           lr_scheduler=lr_scheduler, mpu=mpu, dist_init_required=dist_init_required,
                                           collate_fn=collate_fn, config=cfg_dict, config_class=config_class)
        else:
            engine = DeepSpeedEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                     training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                     dist_init_required=dist_init_required, collate_fn=collate_fn, config=cfg_dict,
                                     config_class=config_class, mesh_device=mesh_device)
    finally:
        zero.restore_init_context()
    return engine, engine.optimizer, engine.training_dataloader, engine.lr_scheduler


def _add_core_arguments(parser):
    group = parser.add_argument_group("DeepSpeed", "DeepSpeed configurations")
    group.add_argument("--deepspeed", default=False, action="store_true",
                       help="Enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepspeed_config", default=None, type=str, help="DeepSpeed json configuration file.")
    group.add_argument("--deepscale", default=False, action="store_true", help=argparse.SUPPRESS)
    group.add_argument("--deepscale_config", default=None, type=str, help=argparse.SUPPRESS)
    return parser


def add_config_arguments(parser):
    """Add ``--deepspeed`` / ``--deepspeed_config`` to an argparse parser."""
    return _add_core_arguments(parser)


def default_inference_config():
    from .inference.config import DeepSpeedInferenceConfig
    return DeepSpeedInferenceConfig().model_dump()


def init_inference(model, config=None, **kwargs):
    """Inference engine (AutoTP / kernel-inject front-end). See :mod:`deepspeed_b200.inference`."""
    from .inference.config import DeepSpeedInferenceConfig
    from .inference.engine import InferenceEngine
    if config is None:
        config = {}
    if isinstance(config, (str, os.PathLike)):
        config = load_ds_config(config)
    elif not isinstance(config, dict):
        raise ValueError(f"'config' argument expected string or dictionary, got {type(config)}")
    overlap = set(config.keys()) & set(kwargs.keys())
    for k in overlap:
        if config[k] != kwargs[k]:
            raise ValueError(f"Conflicting argument '{k}' in 'config':{config[k]} and kwargs:{kwargs[k]}")
    config = dict(config)
    config.update(kwargs)
    return InferenceEngine(model, config=DeepSpeedInferenceConfig(**config))


def tp_model_init(model, tp_size, dtype, config=None, **kwargs):
    """Record tensor-parallel init args to be merged into the ds_config by ``initialize`` (reference :408)."""
    global _TP_MODEL_INIT_ARGS
    if hasattr(model, "ds_autotp_parsed"):
        logger.warning("`tp_model_init` was already called on this model")
        return model
    _TP_MODEL_INIT_ARGS = dict(tp_size=tp_size, dtype=dtype, config=config)
    from .parallel.tp import apply_autotp
    from .config.sections import TPTrainingConfig
    dist.init_distributed()
    groups.initialize(tp_size=tp_size, force=True)
    apply_autotp(model, TPTrainingConfig(autotp_size=tp_size, dtype=dtype), groups._get_model_parallel_group())
    model.ds_autotp_parsed = True
    return model


def merge_tp_model_init_into_config(config_dict, mpu, mesh_param, dist_module):
    """Fold the arguments recorded by :func:`tp_model_init` into the ``tensor_parallel`` section; conflicting values
    raise (reference ``runtime/tensor_parallel/init_utils.py:79``)."""
    if _TP_MODEL_INIT_ARGS is None:
        return
    tp_size, dtype = _TP_MODEL_INIT_ARGS["tp_size"], _TP_MODEL_INIT_ARGS.get("dtype")
    tp_group = _TP_MODEL_INIT_ARGS.get("tp_group")
    if tp_group is not None and mpu is not None:
        raise ValueError("tp_model_init provided tp_group; deepspeed.initialize must not receive mpu.")
    section = config_dict.setdefault("tensor_parallel", {})
    have = section.get("autotp_size")
    if have is not None and have != tp_size:
        raise ValueError(f"Conflicting tensor_parallel.autotp_size in config ({have}) and tp_model_init ({tp_size}).")
    section["autotp_size"] = tp_size
    inner = section.get("tp") or {}
    if not isinstance(inner, dict):
        raise ValueError("tensor_parallel.tp must be a dict when provided.")
    if inner.get("tp_size") not in (None, tp_size):
        raise ValueError(f"Conflicting tensor_parallel.tp.tp_size in config ({inner['tp_size']}) and tp_model_init "
                         f"({tp_size}).")
    inner["tp_size"] = tp_size
    if tp_group is not None:
        if inner.get("tp_group") not in (None, tp_group):
            raise ValueError("Conflicting tensor_parallel.tp.tp_group in config and tp_model_init.")
        inner["tp_group"] = tp_group
    section["tp"] = inner
    if dtype is not None and "dtype" not in section:
        section["dtype"] = str(dtype).replace("torch.", "")


def set_optimizer_flags(config_class, model):
    return None
