"""``deepspeed.utils`` public names (reference ``utils/__init__.py``). Everything except the logger resolves lazily:
``comm`` imports ``utils.logging`` while several of these modules import ``comm``."""
import importlib

from .logging import logger, log_dist, log_dist_once, set_log_level_from_string  # noqa: F401

_LAZY = {
    "get_caller_func": "comms_logging",
    "OnDevice": "init_on_device",
    "instrument_w_nvtx": "nvtx",
    "get_numactl_cmd": "numa",
    "link_hp_params": "mixed_precision_linkage",
    "lazy_init_hp_params_optimizer_state": "mixed_precision_linkage",
}
for _n in ("tensor_fragment", "get_full_hp_param", "get_hp_fragment_mapping", "fragment_address", "get_full_hp_grad",
           "map_to_flat_opt_states", "safe_get_full_fp32_param", "safe_get_full_grad", "safe_get_full_optimizer_state",
           "set_full_hp_param", "set_full_hp_grad", "safe_set_full_fp32_param", "safe_set_full_optimizer_state",
           "safe_set_full_grad", "safe_get_local_fp32_param", "safe_get_local_grad", "safe_get_local_optimizer_state",
           "safe_set_local_fp32_param", "safe_set_local_grad", "safe_set_local_optimizer_state",
           "safe_update_full_grad_vectorized"):
    _LAZY[_n] = "tensor_fragment"
for _n in ("set_z3_leaf_modules", "unset_z3_leaf_modules", "get_z3_leaf_modules", "z3_leaf_module", "z3_leaf_parameter",
           "set_z3_leaf_module", "set_z3_leaf_modules_by_name", "set_z3_leaf_modules_by_suffix"):
    _LAZY[_n] = "z3_leaf_module"
_SUBMODULES = ("groups", "nvtx", "numa", "timer", "memory", "tensor_fragment", "z3_leaf_module", "init_on_device", "bwc",
               "torch", "types", "exceptions", "debug", "comms_logging", "mixed_precision_linkage", "pin_memory",
               "pin_memory_tracker", "static_cache", "config", "zero_to_fp32", "logging")


def __getattr__(name):
    if name == "RepeatingLoader":
        from ..runtime.dataloader import RepeatingLoader
        return RepeatingLoader
    if name in _SUBMODULES:
        return importlib.import_module(f"{__name__}.{name}")
    if name in _LAZY:
        mod = importlib.import_module(f"{__name__}.{_LAZY[name]}")
        if name == "tensor_fragment":  # the reference exports the class under the module's name
            return mod.tensor_fragment
        return getattr(mod, name)
    if not name.startswith("__"):
        groups = importlib.import_module(f"{__name__}.groups")  # reference: ``from .groups import *``
        if hasattr(groups, name):
            return getattr(groups, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
