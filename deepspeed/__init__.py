"""``import deepspeed`` compatibility package: the reference's import paths resolved onto ``deepspeed_b200``.

A user switching from the reference keeps ``import deepspeed``, ``deepspeed.initialize``, ``deepspeed.zero.Init``,
``deepspeed.comm``, ``from deepspeed.ops.adam import FusedAdam`` ... Module paths that moved are translated by
``_ALIASES``; everything else maps ``deepspeed.<path>`` -> ``deepspeed_b200.<path>``. Checkpoints pickle a few small
objects (loss scaler, fragment addresses) under the reference's module names so files are interchangeable with
the reference's tools; this package is what makes those names importable here.
"""
import importlib
import types
import importlib.abc
import importlib.util
import sys

import deepspeed_b200 as _impl

_ALIASES = {
    "deepspeed.runtime.fp16.loss_scaler": "deepspeed_b200.runtime.loss_scaler",
    "deepspeed.runtime.config": "deepspeed_b200.config.config",
    "deepspeed.runtime.config_utils": "deepspeed_b200.config.base",
    "deepspeed.runtime.zero.config": "deepspeed_b200.config.sections",
    "deepspeed.runtime.zero.offload_config": "deepspeed_b200.config.sections",
    "deepspeed.runtime.zero.partition_parameters": "deepspeed_b200.runtime.zero.init_context",
    "deepspeed.runtime.zero.stage3": "deepspeed_b200.runtime.zero.sharded_optimizer",
    "deepspeed.runtime.zero.stage_1_and_2": "deepspeed_b200.runtime.zero.sharded_optimizer",
    "deepspeed.runtime.activation_checkpointing.checkpointing": "deepspeed_b200.runtime.activation_checkpointing",
    "deepspeed.runtime.activation_checkpointing": "deepspeed_b200.runtime",
    "deepspeed.checkpointing": "deepspeed_b200.runtime.activation_checkpointing",
    "deepspeed.pipe": "deepspeed_b200.runtime.pipe",
    "deepspeed.ops.adam": "deepspeed_b200.ops.adam",
    "deepspeed.ops.lamb": "deepspeed_b200.ops.lamb",
    "deepspeed.ops.lion": "deepspeed_b200.ops.lion",
    "deepspeed.ops.adagrad": "deepspeed_b200.ops.adagrad",
    "deepspeed.module_inject": "deepspeed_b200.parallel",
    "deepspeed.module_inject.replace_module": "deepspeed_b200.parallel.replace_module",
    "deepspeed.module_inject.module_quantize": "deepspeed_b200.parallel.replace_module",
    "deepspeed.module_inject.replace_policy": "deepspeed_b200.parallel.replace_module",
    "deepspeed.module_inject.policy": "deepspeed_b200.parallel.replace_module",
    "deepspeed.module_inject.layers": "deepspeed_b200.parallel.layers",
    "deepspeed.module_inject.autotp_config": "deepspeed_b200.parallel.autotp_config",
    "deepspeed.module_inject.auto_tp": "deepspeed_b200.parallel.tp",
    "deepspeed.runtime.model_checkpointing": "deepspeed_b200.runtime.model_checkpointing",
    "deepspeed.runtime.model_checkpointing.constants": "deepspeed_b200.runtime.model_checkpointing.constants",
    "deepspeed.runtime.model_checkpointing.config": "deepspeed_b200.runtime.model_checkpointing.config",
    "deepspeed.runtime.model_checkpointing.utils": "deepspeed_b200.runtime.model_checkpointing.utils",
    "deepspeed.runtime.model_checkpointing.writer_factory": "deepspeed_b200.runtime.model_checkpointing.writer_factory",
    "deepspeed.runtime.model_checkpointing.data_parallel_writer_factory":
    "deepspeed_b200.runtime.model_checkpointing.data_parallel_writer_factory",
    "deepspeed.io.base_file_writer": "deepspeed_b200.io.base_file_writer",
    "deepspeed.io.base_io_buffer": "deepspeed_b200.io.io_buffers",
    "deepspeed.io.single_io_buffer": "deepspeed_b200.io.io_buffers",
    "deepspeed.io.double_io_buffer": "deepspeed_b200.io.io_buffers",
    "deepspeed.runtime.zero.mics": "deepspeed_b200.runtime.zero.mics",
    "deepspeed.runtime.zero.mics_utils": "deepspeed_b200.runtime.zero.mics",
    "deepspeed.comm.backend": "deepspeed_b200.comm.backend",
    "deepspeed.comm.utils": "deepspeed_b200.comm.utils",
    "deepspeed.io.utils": "deepspeed_b200.io.utils",
    "deepspeed.comm.torch": "deepspeed_b200.comm.backend",
    "deepspeed.comm.config": "deepspeed_b200.comm.config",
    "deepspeed.module_inject.fusedqkv_utils": "deepspeed_b200.parallel.fusedqkv_utils",
    "deepspeed.module_inject.tp_shard": "deepspeed_b200.parallel.fusedqkv_utils",
    "deepspeed.module_inject.auto_tp_model_utils": "deepspeed_b200.parallel.auto_tp_model_utils",
    "deepspeed.moe.mappings": "deepspeed_b200.moe.mappings",
    "deepspeed.module_inject.tp_plan_converter": "deepspeed_b200.parallel.tp_plan_converter",
    "deepspeed.runtime.tensor_parallel": "deepspeed_b200.runtime.tensor_parallel",
    "deepspeed.runtime.tensor_parallel.tp_manager": "deepspeed_b200.runtime.tensor_parallel.tp_manager",
    "deepspeed.runtime.tensor_parallel.init_utils": "deepspeed_b200.runtime.tensor_parallel.init_utils",
    "deepspeed.launcher.launcher_helper": "deepspeed_b200.launcher.launcher_helper",
    "deepspeed.elasticity.config": "deepspeed_b200.elasticity.config",
    "deepspeed.module_inject.auto_ep": "deepspeed_b200.parallel.auto_ep",
    "deepspeed.module_inject.auto_ep_layer": "deepspeed_b200.parallel.auto_ep",
    "deepspeed.module_inject.auto_ep_config": "deepspeed_b200.parallel.auto_ep_presets",
    "deepspeed.module_inject.auto_ep_folding": "deepspeed_b200.parallel.auto_ep_folding",
    "deepspeed.module_inject.auto_ep_presets": "deepspeed_b200.parallel.auto_ep_presets",
    "deepspeed.module_inject.auto_ep_presets.base": "deepspeed_b200.parallel.auto_ep_presets",
    "deepspeed.module_inject.auto_ep_presets.registry": "deepspeed_b200.parallel.auto_ep_presets",
    "deepspeed.moe.ep_router": "deepspeed_b200.parallel.auto_ep",
    "deepspeed.moe.ep_experts": "deepspeed_b200.parallel.auto_ep",
    "deepspeed.moe.fused_expert_layout": "deepspeed_b200.moe.fused_expert_layout",
    "deepspeed.model_implementations.transformers.ds_transformer":
    "deepspeed_b200.ops.transformer.inference.ds_transformer",
    "deepspeed.ops.transformer.inference.ds_attention": "deepspeed_b200.ops.transformer.inference.ds_transformer",
    "deepspeed.ops.transformer.inference.ds_mlp": "deepspeed_b200.ops.transformer.inference.ds_transformer",
    "deepspeed.ops.adam.zenflow_cpu_adam": "deepspeed_b200.ops.adam.zenflow",
    "deepspeed.ops.adam.zenflow_torch_adam": "deepspeed_b200.ops.adam.zenflow",
    "deepspeed.runtime.zero.leaf_module_config": "deepspeed_b200.config.sections",
    "deepspeed.runtime.zero.parameter_offload": "deepspeed_b200.runtime.zero.param_status",
    "deepspeed.inference.v2.config_v2": "deepspeed_b200.inference.v2.engine_factory",
    "deepspeed.checkpoint.reshape_utils": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.reshape_3d_utils": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.reshape_meg_2d": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.zero_checkpoint": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.deepspeed_checkpoint": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.utils": "deepspeed_b200.checkpoint.reshape",
    "deepspeed.checkpoint.universal_checkpoint": "deepspeed_b200.checkpoint.universal",
    "deepspeed.nvme.perf_run_sweep": "deepspeed_b200.nvme",
    "deepspeed.nvme.perf_generate_param": "deepspeed_b200.nvme",
    "deepspeed.nvme.test_ds_aio": "deepspeed_b200.nvme",
    "deepspeed.accelerator.real_accelerator": "deepspeed_b200.accelerator",
    "deepspeed.accelerator.abstract_accelerator": "deepspeed_b200.accelerator",
    "deepspeed.git_version_info": "deepspeed_b200.version",
    "deepspeed.utils.zero_to_fp32": "deepspeed_b200.checkpoint.zero_to_fp32",
    "deepspeed.checkpoint.ds_to_universal": "deepspeed_b200.checkpoint.ds_to_universal",
    "deepspeed.sequence.layer": "deepspeed_b200.sequence.layer",
    "deepspeed.runtime.sequence_parallel.ulysses_sp": "deepspeed_b200.sequence.ulysses_sp",
    "deepspeed.runtime.sequence_parallel": "deepspeed_b200.sequence",
}


class _AliasModule(types.ModuleType):
    """Stand-in for an aliased module: attribute reads go to the real module. It is a separate object so that the
    import system, which binds every imported submodule as an attribute of its parent, writes those bindings here
    and not into the real package (an alias whose leaf name equals a real attribute with a different target -
    ``deepspeed.comm.torch``, ``deepspeed.runtime.activation_checkpointing`` - would otherwise overwrite it)."""

    def __init__(self, name, real):
        super().__init__(name, getattr(real, "__doc__", None))
        object.__setattr__(self, "_ds_real", real)
        if hasattr(real, "__path__"):
            self.__dict__["__path__"] = []
        self.__dict__["__all__"] = list(getattr(real, "__all__", None) or
                                        [k for k in vars(real) if not k.startswith("_")])

    def __getattr__(self, name):
        return getattr(object.__getattribute__(self, "_ds_real"), name)

    def __setattr__(self, name, value):
        if isinstance(value, types.ModuleType) or name.startswith("__"):
            self.__dict__[name] = value  # submodule binding made by the import system
        else:
            setattr(object.__getattribute__(self, "_ds_real"), name, value)  # a user patching a global

    def __dir__(self):
        return sorted(set(dir(object.__getattribute__(self, "_ds_real"))) | set(self.__dict__))


class _AliasLoader(importlib.abc.Loader):

    def __init__(self, target):
        self.target = target

    def create_module(self, spec):
        return _AliasModule(spec.name, importlib.import_module(self.target))

    def exec_module(self, module):
        pass


class _AliasFinder(importlib.abc.MetaPathFinder):

    def find_spec(self, fullname, path, target=None):
        if not fullname.startswith("deepspeed."):
            return None
        tgt = _ALIASES.get(fullname, "deepspeed_b200." + fullname[len("deepspeed."):])
        try:
            importlib.import_module(tgt)
        except ImportError:
            return None
        spec = importlib.util.spec_from_loader(fullname, _AliasLoader(tgt))
        return spec


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())

from deepspeed_b200 import *  # noqa: E402,F401,F403
from deepspeed_b200 import (__version__, add_config_arguments, comm, init_distributed, init_inference,  # noqa: E402
                            initialize, tp_model_init, zero, DeepSpeedEngine, DeepSpeedConfig)
from deepspeed_b200 import comm as dist  # noqa: E402,F401


def __getattr__(name):
    return getattr(_impl, name)
