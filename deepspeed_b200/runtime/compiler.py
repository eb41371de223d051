"""``torch.compile`` helpers (reference ``runtime/compiler.py``)."""
import torch

from ..utils.torch import required_torch_version


def is_compile_supported():
    return required_torch_version(min_version=2.1) and hasattr(torch, "compile")


def disable(func):
    """Keep ``func`` out of traced graphs (collective launchers, hooks)."""
    return torch.compiler.disable(func) if is_compile_supported() else func


def enable(min_version=None):

    def deco(func):
        return func

    return deco


def is_compiling():
    c = getattr(torch, "compiler", None)
    return bool(c is not None and hasattr(c, "is_compiling") and c.is_compiling())
