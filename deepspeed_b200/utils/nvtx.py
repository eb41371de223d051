"""NVTX range decorator (reference ``utils/nvtx.py:9-38``): ranges named by ``func.__qualname__``."""
import functools

import torch

_ENABLED = True


def enable_nvtx(flag=True):
    global _ENABLED
    _ENABLED = flag


def instrument_w_nvtx(func):
    if not torch.cuda.is_available():
        return func

    @functools.wraps(func)
    def wrapped(*args, **kwargs):
        if not _ENABLED:
            return func(*args, **kwargs)
        torch.cuda.nvtx.range_push(func.__qualname__)
        try:
            return func(*args, **kwargs)
        finally:
            torch.cuda.nvtx.range_pop()

    return wrapped
