"""CUDA-graph replay of the decode forward during HybridEngine generation.

Reference: ``runtime/hybrid_engine_graph.py`` (``DecodeGraphCache``, ``decode_steps_from_generate_kwargs``,
``validate_cuda_graph_support``). A decode step is a few hundred to a thousand small kernels: on a B200 they finish
before the host has queued the next step, so generation is launch-bound and one graph replay per token removes that.

Graphs are keyed by the *signature* of a decode call - the shapes / dtypes of its tensor arguments and the identity of
its non-tensor ones (the KV-cache object). A model that reads the write position from a device tensor (static cache +
``cache_position``) therefore runs every position of a sequence from ONE graph; one whose attention mask grows by a
column per token gets one graph per position, which is what the reference does for every model (its kernels take the
position as a launch parameter). ``max_positions`` bounds the number of graphs kept. A sequence runs entirely on
graphs or entirely eagerly (``begin_sequence``): mixing the two is only safe for models without host-side position
state, and nothing is gained by it.

The graph backend is injectable so that the dispatch logic is testable without a GPU."""
import torch

from ..utils.logging import logger


class _CudaGraphBackend:

    def capture(self, fn, pool):
        """Warm up on a side stream (lazy allocations, autotuning), then record ``fn``; returns (graph, output, pool)."""
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fn()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, pool=pool):
            out = fn()
        return graph, out, graph.pool()

    def replay(self, graph):
        graph.replay()


def _tensor_names(kwargs):
    return [k for k, v in kwargs.items() if torch.is_tensor(v)]


def _signature(args, kwargs):
    sig = [("*", tuple(a.shape), a.dtype) if torch.is_tensor(a) else ("*", id(a)) for a in args]
    for k in sorted(kwargs):
        v = kwargs[k]
        if torch.is_tensor(v):
            sig.append((k, tuple(v.shape), v.dtype))
        elif isinstance(v, (bool, int, float, str, type(None))):
            sig.append((k, v))
        else:
            sig.append((k, id(v)))
    return tuple(sig)


class DecodeGraphCache:
    """Callable that stands in for the module's ``forward`` while generating."""

    def __init__(self, forward, max_positions, backend=None):
        self._forward = forward
        self._max_positions = max_positions
        self._backend = backend if backend is not None else _CudaGraphBackend()
        self._graphs = {}  # signature -> (graph, static args, static kwargs, output)
        self._pool = None
        self._position = 0
        self._declared = None
        self._use_graphs = False
        self._disabled = False
        self.replays = 0

    @property
    def captured_positions(self):
        return len(self._graphs)

    def invalidate(self):
        self._graphs.clear()
        self._pool = None

    def begin_sequence(self, num_decode_steps):
        """Decide before the first decode step whether this sequence replays graphs. ``None`` / non-positive (length
        not pinned) or more steps than ``max_positions``: eager."""
        self._position = 0
        self._declared = num_decode_steps
        self._use_graphs = bool(not self._disabled and num_decode_steps is not None
                                and 0 < num_decode_steps <= self._max_positions)

    def _fall_back(self, reason, args, kwargs):
        logger.warning(f"HybridEngine CUDA graph disabled: {reason}")
        self.invalidate()
        self._disabled = True
        self._use_graphs = False
        return self._forward(*args, **kwargs)

    def _capture(self, sig, args, kwargs):
        if len(self._graphs) >= self._max_positions:  # bounded memory: drop the oldest graph
            self._graphs.pop(next(iter(self._graphs)))
        s_args = tuple(a.clone() if torch.is_tensor(a) else a for a in args)
        s_kwargs = dict(kwargs)
        for k in _tensor_names(kwargs):
            s_kwargs[k] = kwargs[k].clone()
        graph, out, self._pool = self._backend.capture(lambda: self._forward(*s_args, **s_kwargs), self._pool)
        self._graphs[sig] = (graph, s_args, s_kwargs, out)

    def __call__(self, *args, **kwargs):
        ids = kwargs.get("input_ids", args[0] if args and torch.is_tensor(args[0]) else None)
        decode = ids is not None and ids.dim() == 2 and ids.shape[1] == 1
        if not self._use_graphs or not decode:
            return self._forward(*args, **kwargs)
        position = self._position
        self._position += 1
        if position >= self._declared:
            raise RuntimeError(f"HybridEngine CUDA graph: generation exceeded the declared length ({self._declared} "
                               "decode steps). Pin the length (min_new_tokens == max_new_tokens) within "
                               "hybrid_engine.max_out_tokens, or disable hybrid_engine.enable_cuda_graph.")
        sig = _signature(args, kwargs)
        if sig not in self._graphs:
            try:
                self._capture(sig, args, kwargs)
            except Exception as e:  # noqa: BLE001
                if self.replays and position > 0:
                    raise
                return self._fall_back(f"capture failed at position {position}: {e}", args, kwargs)
        graph, s_args, s_kwargs, out = self._graphs[sig]
        for a, s in zip(args, s_args):
            if torch.is_tensor(a) and a.data_ptr() != s.data_ptr():
                s.copy_(a)
        for k in _tensor_names(kwargs):
            if kwargs[k].data_ptr() != s_kwargs[k].data_ptr():
                s_kwargs[k].copy_(kwargs[k])
        self._backend.replay(graph)
        self.replays += 1
        return out


def decode_steps_from_generate_kwargs(kwargs):
    """Single-token forwards of a ``generate`` call when its length is pinned (``min_new_tokens == max_new_tokens``):
    one prompt forward yields the first token, every further token is one decode step. ``None`` otherwise - a
    sequence that may stop early or run long cannot commit to graphs up front."""
    max_new, min_new = kwargs.get("max_new_tokens"), kwargs.get("min_new_tokens")
    if max_new is None or min_new is None or max_new != min_new:
        return None
    return int(max_new) - 1


def validate_cuda_graph_support(config, zero_stage):
    """Reason why decode graphs cannot be used, or ``None``."""
    if not torch.cuda.is_available():
        return "CUDA graphs require a CUDA device"
    if zero_stage == 3:
        # parameters are gathered into fresh buffers for each generate call; a graph would keep reading the old ones
        return "CUDA graphs are not supported with ZeRO stage 3"
    if getattr(config, "release_inference_cache", False):
        return "CUDA graphs are not supported with release_inference_cache"
    if getattr(config, "inference_tp_size", 1) > 1:
        return "CUDA graphs are not supported with inference_tp_size > 1"
    return None
