"""``see_memory_usage`` (reference ``runtime/utils.py:815``) and small memory helpers."""
import gc

import psutil
import torch

from .logging import logger


def see_memory_usage(message, force=False, ranks=(0, )):
    if not force:
        return
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_rank() not in ranks:
        return
    gc.collect()
    logger.info(message)
    if torch.cuda.is_available():
        gb = 1024**3
        logger.info(f"MA {torch.cuda.memory_allocated() / gb:.2f} GB  Max_MA {torch.cuda.max_memory_allocated() / gb:.2f} GB"
                    f"  CA {torch.cuda.memory_reserved() / gb:.2f} GB  Max_CA {torch.cuda.max_memory_reserved() / gb:.2f} GB")
        torch.cuda.reset_peak_memory_stats()
    vm = psutil.virtual_memory()
    logger.info(f"CPU Virtual Memory:  used = {(vm.total - vm.available) / 1024**3:.2f} GB, percent = {vm.percent}%")


def empty_cache():
    if torch.cuda.is_available():
        torch.cuda.empty_cache()
