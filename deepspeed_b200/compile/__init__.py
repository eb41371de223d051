"""DeepCompile counterpart.

The reference's DeepCompile (``deepspeed/compile``: ``backend.py:284 make_backend``, ``init_z3.py``, FX passes for
all-gather / release / prefetch / offload insertion, the native executor ``csrc/compile/*.cpp``) exists because its
eager ZeRO-3 pays a Python hook per ``nn.Module`` and cannot reorder collectives without rewriting the graph. This
framework schedules ZeRO at *unit* granularity with explicit CUDA streams (``runtime/zero/coordinator.py``: one-unit-ahead
prefetch, reduce-scatter on a side stream) and launch-bound inner loops are captured as CUDA graphs
(:class:`CUDAGraphRunner`) - no tracing compiler sits on the hot path, by design.

What is provided: the config model, ``is_deepcompile_supported`` / ``init_z3`` entry points that configure the
equivalent eager mechanisms (prefetch depth, persistence, selective unshard by memory budget) and report what they
did, ``make_backend`` returning an eager backend for ``torch.compile`` callers, and the graph runner."""
from .config import CompileConfig  # noqa: F401
from .backend import make_backend, init_z1, init_z3, is_deepcompile_supported  # noqa: F401
from .graph_capture import CUDAGraphRunner  # noqa: F401
