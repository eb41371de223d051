"""Native pinned host memory (reference ``utils/pin_memory.py`` ``NativePinnedMemory``): page-aligned, ``mlock``-ed
allocations from the host runtime (``csrc/cpu/pin_memory.cpp``), registered with CUDA when a device is present so
copies from them are true asynchronous DMA. Selected with ``DS_PIN_MEMORY_BACKEND=native``."""
import ctypes
import os
import weakref

import torch

from .pin_memory_tracker import track_pinned_memory


class NativePinnedMemory:

    def __init__(self):
        from ..ops import cpu_lib
        self._lib = cpu_lib.lib()
        if not hasattr(self._lib, "dsb200_locked_alloc"):
            raise RuntimeError("DS_PIN_MEMORY_BACKEND=native requires the host runtime library (libdsb200_cpu.so)")
        self._ranges = {}  # base address -> (end address, cuda registered)
        self._finalizers = {}

    def _new_locked(self, nbytes):
        locked = ctypes.c_int(0)
        ptr = self._lib.dsb200_locked_alloc(max(int(nbytes), 1), ctypes.byref(locked))
        if not ptr:
            raise MemoryError(f"could not allocate {nbytes} page-aligned host bytes")
        registered = False
        if torch.cuda.is_available():
            try:  # cudaHostRegisterDefault: the pages become DMA-able for cudaMemcpyAsync
                registered = int(torch.cuda.cudart().cudaHostRegister(ptr, max(int(nbytes), 1), 0)) == 0
            except Exception:  # noqa: BLE001
                registered = False
        return ptr, registered

    def pin(self, tensor, make_copy=True, match_shape=True):
        numel, esz = tensor.numel(), tensor.element_size()
        nbytes = numel * esz
        ptr, registered = self._new_locked(nbytes)
        raw = (ctypes.c_uint8 * max(nbytes, 1)).from_address(ptr)
        base = torch.frombuffer(raw, dtype=torch.uint8, count=max(nbytes, 1))
        out = base[:nbytes].view(tensor.dtype) if nbytes else torch.empty(0, dtype=tensor.dtype)
        if make_copy and nbytes:
            out.copy_(tensor.detach().reshape(-1).to("cpu"))
        if match_shape:
            out = out.view(tensor.shape)
        self._ranges[ptr] = (ptr + max(nbytes, 1), registered)
        out.ds_pinned = True
        out.ds_pin_base = ptr
        track_pinned_memory(nbytes)
        # free when the storage dies, unless unpin() did it first. torch.frombuffer keeps ``raw`` referenced for as long
        # as the storage (i.e. any view of it) lives, so ``raw`` is the object whose collection means "no users left"
        # (the Python wrapper of ``base`` can be collected long before its views).
        self._finalizers[ptr] = weakref.finalize(raw, NativePinnedMemory._release, self._lib, ptr, registered,
                                                 self._ranges, self._finalizers)
        return out

    def is_pinned(self, tensor):
        if getattr(tensor, "ds_pinned", False):
            return True
        if tensor.device.type != "cpu" or tensor.numel() == 0:
            return False
        p = tensor.data_ptr()
        return any(b <= p < e for b, (e, _) in self._ranges.items())

    def unpin(self, tensor):
        """Free the allocation behind ``tensor``; the tensor (and its views) must not be used afterwards."""
        begin = getattr(tensor, "ds_pin_base", None)
        if begin is None:
            begin = tensor.data_ptr()
        fin = self._finalizers.pop(begin, None)
        if fin is not None:
            fin.detach()
        ent = self._ranges.pop(begin, None)
        if ent is None:
            return False
        if hasattr(tensor, "ds_pinned"):
            tensor.ds_pinned = False
        return NativePinnedMemory._free(self._lib, begin, ent[1])

    @staticmethod
    def _free(lib, ptr, registered):
        if registered:
            try:
                torch.cuda.cudart().cudaHostUnregister(ptr)
            except Exception:  # noqa: BLE001
                pass
        return bool(lib.dsb200_locked_free(ptr))

    @staticmethod
    def _release(lib, ptr, registered, ranges, finalizers):
        ranges.pop(ptr, None)
        finalizers.pop(ptr, None)
        try:
            NativePinnedMemory._free(lib, ptr, registered)
        except Exception:  # noqa: BLE001  (interpreter shutdown)
            pass


_SHARED = None


def get_native_pinned_memory():
    global _SHARED
    if _SHARED is None:
        _SHARED = NativePinnedMemory()
    return _SHARED


def get_active_native_pinned_memory():
    """The shared manager when ``DS_PIN_MEMORY_BACKEND=native``, else None (torch pinning is used)."""
    if os.environ.get("DS_PIN_MEMORY_BACKEND", "torch").strip().lower() != "native":
        return None
    return get_native_pinned_memory()
