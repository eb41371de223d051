"""``get_accelerator()`` compatibility shim. The reference's 72-method multi-backend abstraction
(``accelerator/abstract_accelerator.py``, 9 backends) is dropped by design: there is one device family (CUDA/B200,
with CPU for the plumbing tests). User code that calls ``get_accelerator().xxx`` keeps working."""
import os

import torch


class B200Accelerator:
    _name = "cuda"

    def is_synchronized_device(self):
        return False

    def use_host_timers(self):
        return not torch.cuda.is_available()

    def resolves_data_dependency(self):
        return False

    def handles_memory_backpressure(self):
        return False

    def device_name(self, device_index=None):
        base = "cuda" if torch.cuda.is_available() else "cpu"
        return base if device_index is None or base == "cpu" else f"cuda:{device_index}"

    def device(self, device_index=None):
        return torch.device(self.device_name(device_index))

    def set_device(self, device_index):
        if torch.cuda.is_available():
            torch.cuda.set_device(device_index)

    def current_device(self):
        return torch.cuda.current_device() if torch.cuda.is_available() else 0

    def current_device_name(self):
        return self.device_name(self.current_device())

    def device_count(self):
        return torch.cuda.device_count() if torch.cuda.is_available() else 1

    def synchronize(self, device_index=None):
        if torch.cuda.is_available():
            torch.cuda.synchronize(device_index)

    def is_available(self):
        return torch.cuda.is_available()

    def communication_backend_name(self):
        return "nccl" if torch.cuda.is_available() else "gloo"

    def manual_seed(self, seed):
        torch.manual_seed(seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed(seed)

    def manual_seed_all(self, seed):
        torch.manual_seed(seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(seed)

    def get_rng_state(self, device_index=None):
        return torch.cuda.get_rng_state(device_index) if torch.cuda.is_available() else torch.get_rng_state()

    def set_rng_state(self, new_state, device_index=None):
        if torch.cuda.is_available():
            torch.cuda.set_rng_state(new_state, device_index if device_index is not None else "cuda")
        else:
            torch.set_rng_state(new_state)

    def Stream(self, *a, **k):
        return torch.cuda.Stream(*a, **k) if torch.cuda.is_available() else None

    def stream(self, s):
        import contextlib
        return torch.cuda.stream(s) if torch.cuda.is_available() else contextlib.nullcontext()

    def current_stream(self, device_index=None):
        return torch.cuda.current_stream(device_index) if torch.cuda.is_available() else None

    def default_stream(self, device_index=None):
        return torch.cuda.default_stream(device_index) if torch.cuda.is_available() else None

    def Event(self, **kw):
        return torch.cuda.Event(**kw) if torch.cuda.is_available() else None

    def empty_cache(self):
        if torch.cuda.is_available():
            torch.cuda.empty_cache()

    def memory_allocated(self, device_index=None):
        return torch.cuda.memory_allocated(device_index) if torch.cuda.is_available() else 0

    def max_memory_allocated(self, device_index=None):
        return torch.cuda.max_memory_allocated(device_index) if torch.cuda.is_available() else 0

    def reset_peak_memory_stats(self, device_index=None):
        if torch.cuda.is_available():
            torch.cuda.reset_peak_memory_stats(device_index)

    def memory_reserved(self, device_index=None):
        return torch.cuda.memory_reserved(device_index) if torch.cuda.is_available() else 0

    def total_memory(self, device_index=None):
        return torch.cuda.get_device_properties(device_index or 0).total_memory if torch.cuda.is_available() else 0

    def available_memory(self, device_index=None):
        return torch.cuda.mem_get_info(device_index)[0] if torch.cuda.is_available() else 0

    def is_bf16_supported(self):
        return True

    def is_fp16_supported(self):
        return True

    def supported_dtypes(self):
        return [torch.float, torch.half, torch.bfloat16]

    def is_triton_supported(self):
        return False  # no Triton on hot paths by design

    def pin_memory(self, tensor, align_bytes=1):
        return tensor.pin_memory() if torch.cuda.is_available() else tensor

    def is_pinned(self, tensor):
        return tensor.is_pinned()

    def on_accelerator(self, tensor):
        return tensor.is_cuda

    def range_push(self, msg):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_push(msg)

    def range_pop(self):
        if torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()

    def lazy_call(self, callback):
        callback()

    def op_builder_dir(self):
        return "deepspeed_b200.ops"

    def create_op_builder(self, class_name):
        b = self.get_op_builder(class_name)
        return b() if b is not None else None

    def get_op_builder(self, class_name):
        from ..ops import op_builders
        return getattr(op_builders, class_name, None)

    def build_extension(self):
        from torch.utils.cpp_extension import BuildExtension
        return BuildExtension

    def export_envs(self):
        return ["NCCL", "CUDA", "DSB200"]

    def visible_devices_envs(self):
        return ["CUDA_VISIBLE_DEVICES"]

    def get_compile_backend(self):
        return getattr(self, "_compile_backend", "inductor")

    def set_compile_backend(self, backend):
        import torch._dynamo as dynamo
        if backend not in dynamo.list_backends(exclude_tags=()):
            raise ValueError(f"{backend} not supported by {self.device_name()}. Supported Backends are "
                             f"{dynamo.list_backends(exclude_tags=())}")
        self._compile_backend = backend

    def set_visible_devices_envs(self, current_env, local_accelerator_ids):
        for env in self.visible_devices_envs():
            current_env[env] = ",".join(map(str, local_accelerator_ids))

    # ---- RNG ---------------------------------------------------------------------------------
    @property
    def random(self):
        return torch.cuda.random if torch.cuda.is_available() else torch.random

    def initial_seed(self):
        return torch.cuda.initial_seed() if torch.cuda.is_available() else torch.initial_seed()

    def default_generator(self, device_index):
        return torch.cuda.default_generators[device_index] if torch.cuda.is_available() else torch.default_generator

    # ---- memory statistics (older spellings included) -------------------------------------------
    def _cuda_stat(self, fn, device_index=None, default=0):
        return getattr(torch.cuda, fn)(device_index) if torch.cuda.is_available() else default

    def reset_max_memory_allocated(self, device_index=None):
        if torch.cuda.is_available():
            torch.cuda.reset_peak_memory_stats(device_index)

    def memory_cached(self, device_index=None):
        return self._cuda_stat("memory_reserved", device_index)

    def max_memory_cached(self, device_index=None):
        return self._cuda_stat("max_memory_reserved", device_index)

    def reset_max_memory_cached(self, device_index=None):
        self.reset_max_memory_allocated(device_index)

    def max_memory_reserved(self, device_index=None):
        return self._cuda_stat("max_memory_reserved", device_index)

    def memory_stats(self, device_index=None):
        return torch.cuda.memory_stats(device_index) if torch.cuda.is_available() else {}

    # ---- CUDA graphs ---------------------------------------------------------------------------
    def create_graph(self):
        return torch.cuda.CUDAGraph()

    def capture_to_graph(self, graph, pool=None, stream=None):
        return torch.cuda.graph(graph, pool=pool, stream=stream)

    def replay_graph(self, graph):
        graph.replay()

    def prefer_triton_grouped_mm(self):
        return False  # the grouped GEMM is the tcgen05 kernel (ops/grouped_gemm.py)

    # ---- typed tensor constructors on the device -----------------------------------------------
    def _typed(self, dtype):
        from functools import partial
        return partial(torch.tensor, dtype=dtype, device=self.current_device_name())

    BFloat16Tensor = property(lambda self: self._typed(torch.bfloat16))
    ByteTensor = property(lambda self: self._typed(torch.uint8))
    DoubleTensor = property(lambda self: self._typed(torch.double))
    FloatTensor = property(lambda self: self._typed(torch.float))
    HalfTensor = property(lambda self: self._typed(torch.half))
    IntTensor = property(lambda self: self._typed(torch.int))
    LongTensor = property(lambda self: self._typed(torch.long))

    def unpin_memory(self, tensor):
        """Release a tensor pinned through the native backend (``DS_PIN_MEMORY_BACKEND=native``); torch-pinned tensors
        are freed by their storage."""
        from ..utils.pin_memory import get_active_native_pinned_memory
        mgr = get_active_native_pinned_memory()
        return mgr.unpin(tensor) if mgr is not None and mgr.is_pinned(tensor) else False


DeepSpeedAccelerator = B200Accelerator  # the reference's abstract base; one concrete device family here

_ACC = None


def is_current_accelerator_supported():
    return True


def get_accelerator():
    global _ACC
    if _ACC is None:
        _ACC = B200Accelerator()
    return _ACC


def set_accelerator(acc):
    global _ACC
    _ACC = acc
