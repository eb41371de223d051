"""Pinned (page-locked) host memory manager (reference ``csrc/pin_memory`` / ``op_builder/pin_memory.py``): allocates
page-aligned host tensors, tracks them, and frees them explicitly - used by the NVMe swappers and offload paths."""
import torch


class PinnedTensorManager:

    def __init__(self):
        self._live = {}

    def alloc(self, numel, dtype=torch.uint8):
        t = torch.empty(numel, dtype=dtype, device="cpu", pin_memory=torch.cuda.is_available())
        self._live[t.data_ptr()] = t
        return t

    def free(self, tensor):
        return self._live.pop(tensor.data_ptr(), None) is not None

    def is_managed(self, tensor):
        return tensor.data_ptr() in self._live

    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in self._live.values())


_MANAGER = PinnedTensorManager()


def pin_memory_alloc(numel, dtype=torch.uint8):
    return _MANAGER.alloc(numel, dtype)


def pin_memory_free(tensor):
    return _MANAGER.free(tensor)


class PinMemoryBuilder:
    NAME = "pin_memory"

    def is_compatible(self, verbose=False):
        return True

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
