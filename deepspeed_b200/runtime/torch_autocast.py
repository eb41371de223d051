"""torch.autocast integration (reference ``runtime/torch_autocast.py:35-94``): parameters of "lower precision safe"
modules are tagged with the autocast dtype as their communication dtype, so ZeRO reduces those gradients in bf16/fp16
while the rest stays in fp32; ``autocast_if_enabled`` wraps the engine's forward."""
import importlib
from contextlib import contextmanager
from typing import Iterable, List, Set, Union

import torch

LOWER_PRECISION_SAFE_MODULES = [torch.nn.Linear, torch.nn.Conv1d, torch.nn.Conv2d, torch.nn.Conv3d]
PARAM_COMM_DTYPE_ATTR_NAME = "comm_dtype"
TORCH_AUTOCAST_INITIALIZED = False
TORCH_AUTOCAST_DTYPE = None
_WARNED_NESTED_AUTOCAST = False


def init_autocast_params(engine, dtype: torch.dtype, torch_autocast_lower_precision_safe_modules: Union[None, List[str]]):
    assert not engine.zero_quantized_weights(), "Cannot enable both torch autocast and zero quantized weights"
    if torch_autocast_lower_precision_safe_modules is None:
        classes = list(LOWER_PRECISION_SAFE_MODULES)
    else:
        classes = []
        for name in torch_autocast_lower_precision_safe_modules:
            try:
                pkg, cls = name.rsplit(".", 1)
                classes.append(getattr(importlib.import_module(pkg), cls))
            except Exception as e:
                raise ValueError(f"Failed to import lower precision safe module {name}: {e}")
    for module in engine.module.modules():
        if module.__class__ in classes:
            for p in module.parameters(recurse=False):
                setattr(p, PARAM_COMM_DTYPE_ATTR_NAME, dtype)
    global TORCH_AUTOCAST_INITIALIZED, TORCH_AUTOCAST_DTYPE
    TORCH_AUTOCAST_INITIALIZED, TORCH_AUTOCAST_DTYPE = True, dtype


def is_autocast_initialized() -> bool:
    return TORCH_AUTOCAST_INITIALIZED


def get_default_autocast_lower_precision_modules() -> List[str]:
    return [f"{c.__module__}.{c.__name__}" for c in LOWER_PRECISION_SAFE_MODULES]


def get_autocast_dtype() -> torch.dtype:
    return TORCH_AUTOCAST_DTYPE


def has_comm_dtype(param) -> bool:
    return hasattr(param, PARAM_COMM_DTYPE_ATTR_NAME)


def get_comm_dtype(param) -> torch.dtype:
    return getattr(param, PARAM_COMM_DTYPE_ATTR_NAME, param.dtype)


def get_all_comm_dtypes(params: Iterable) -> Set[torch.dtype]:
    return {get_comm_dtype(p) for p in params}


def sort_dtypes(dtypes: List[torch.dtype]) -> List[torch.dtype]:
    return sorted(dtypes, key=str)


@contextmanager
def autocast_if_enabled(engine):
    """Run the forward under ``torch.autocast`` when the config enables it; warn once if the user nests their own."""
    global _WARNED_NESTED_AUTOCAST
    cfg = engine._config.torch_autocast_config
    if not cfg.enabled:
        yield
        return
    if torch.is_autocast_enabled() and not _WARNED_NESTED_AUTOCAST:
        from ..utils.logging import logger
        logger.warning("torch.autocast is already enabled outside the engine; the engine's setting takes precedence")
        _WARNED_NESTED_AUTOCAST = True
    dtype = getattr(torch, str(cfg.dtype).replace("torch.", ""))
    with torch.autocast(device_type=engine.device.type, dtype=dtype, enabled=True):
        yield
