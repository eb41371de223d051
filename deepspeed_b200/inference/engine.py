"""``InferenceEngine`` (v1): dtype conversion + AutoTP sharding + (optional) CUDA-graph replay of the forward.

Reference: ``inference/engine.py:43`` (``InferenceEngine``: ``_apply_injection_policy`` -> ``replace_module``, TP group
creation, ``_create_cuda_graph`` / ``_graph_replay``, ``forward``, ``generate`` passthrough with checks). Kernel
injection of the reference's per-architecture containers is replaced by the fused ops of this framework for the
model families in :mod:`deepspeed_b200.models`; other HF models run with AutoTP-sharded linears."""
import torch
import torch.nn as nn

from .. import comm as dist
from ..utils import groups
from ..utils.logging import log_dist
from .config import DeepSpeedInferenceConfig


class InferenceEngine(nn.Module):

    def __init__(self, model, config: DeepSpeedInferenceConfig):
        super().__init__()
        self._config = config
        self.module = model
        self.dtype = config.dtype
        self.mp_world_size = config.tensor_parallel.tp_size
        self.mpu = config.mpu
        self.cuda_graph_created = False
        self._graphs = {}
        self.device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else \
            torch.device("cpu")
        if self.mp_world_size > 1:
            dist.init_distributed()
            groups.initialize(tp_size=self.mp_world_size, force=True)
            self.mp_group = groups._get_model_parallel_group()
            from ..parallel.tp import apply_autotp
            from ..config.sections import TPTrainingConfig
            apply_autotp(self.module, TPTrainingConfig(autotp_size=self.mp_world_size), self.mp_group)
        else:
            self.mp_group = None
        if self.dtype in (torch.float16, torch.bfloat16, torch.float32):
            self.module.to(self.dtype)
        if not config.keep_module_on_host:
            self.module.to(self.device)
        self.module.eval()
        if config.checkpoint is not None:
            self._load_checkpoint(config.checkpoint)
        log_dist(f"InferenceEngine: dtype={self.dtype} tp={self.mp_world_size} cuda_graph={config.enable_cuda_graph}", [0])

    def _load_checkpoint(self, ckpt):
        path = ckpt if isinstance(ckpt, str) else ckpt.get("checkpoints", [None])[0]
        if path:
            sd = torch.load(path, map_location="cpu", weights_only=False)
            sd = sd.get("module", sd)
            self.module.load_state_dict(sd, strict=False)

    # ---- CUDA graph of the forward (fixed shapes) ---------------------------------------------------
    def _graph_key(self, args, kwargs):
        return tuple((tuple(a.shape), a.dtype) for a in list(args) + list(kwargs.values()) if torch.is_tensor(a))

    def _create_cuda_graph(self, key, *inputs, **kwargs):
        static_in = [a.clone() if torch.is_tensor(a) else a for a in inputs]
        static_kw = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in kwargs.items()}
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                self.module(*static_in, **static_kw)
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = self.module(*static_in, **static_kw)
        self._graphs[key] = (g, static_in, static_kw, out)
        self.cuda_graph_created = True

    def _graph_replay(self, key, *inputs, **kwargs):
        g, static_in, static_kw, out = self._graphs[key]
        for s, a in zip(static_in, inputs):
            if torch.is_tensor(a):
                s.copy_(a)
        for k, v in kwargs.items():
            if torch.is_tensor(v):
                static_kw[k].copy_(v)
        g.replay()
        return out

    @torch.no_grad()
    def forward(self, *inputs, **kwargs):
        if self._config.enable_cuda_graph and torch.cuda.is_available():
            key = self._graph_key(inputs, kwargs)
            if key not in self._graphs:
                self._create_cuda_graph(key, *inputs, **kwargs)
            return self._graph_replay(key, *inputs, **kwargs)
        return self.module(*inputs, **kwargs)

    @torch.no_grad()
    def generate(self, *inputs, **kwargs):
        if hasattr(self.module, "generate"):
            nb = kwargs.get("num_beams", 1)
            if nb > 1:
                raise NotImplementedError("DeepSpeed does not support `num_beams` > 1")
            return self.module.generate(*inputs, **kwargs)
        from .v2.engine_v2 import greedy_generate
        return greedy_generate(self.module, *inputs, **kwargs)

    # ---- latency profiling (reference engine.py:197, :564) ------------------------------------------------
    def profile_model_time(self, use_cuda_events=True):
        """Record the latency of every forward; read (and clear) with :meth:`model_times`."""
        if not getattr(self, "model_profile_enabled", False):
            self.module.register_forward_pre_hook(self._pre_forward_hook)
            self.module.register_forward_hook(self._post_forward_hook)
        self.model_profile_enabled = True
        self.use_cuda_events = bool(use_cuda_events) and torch.cuda.is_available()
        self._model_times = []

    def _pre_forward_hook(self, module, *inputs, **kwargs):
        if self.use_cuda_events:
            self._ev0 = torch.cuda.Event(enable_timing=True)
            self._ev0.record()
        else:
            import time
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            self._start = time.time()

    def _post_forward_hook(self, module, input, output):
        if self.use_cuda_events:
            ev1 = torch.cuda.Event(enable_timing=True)
            ev1.record()
            ev1.synchronize()
            ms = self._ev0.elapsed_time(ev1)
        else:
            import time
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            ms = (time.time() - self._start) * 1e3
        self._model_times.append(ms)

    def model_times(self):
        assert getattr(self, "model_profile_enabled", False), "model profiling is not enabled"
        times, self._model_times = self._model_times, []
        return times

    # ---- torch.compile passthrough, teardown ----------------------------------------------------------------
    def compile(self, backend=None, compile_kwargs=None):
        if getattr(self, "_is_compiled", False):
            return
        self.module.compile(backend=backend or "inductor", **(compile_kwargs or {}))
        self._is_compiled = True

    @property
    def is_compiled(self):
        return getattr(self, "_is_compiled", False)

    def destroy(self):
        """Drop captured graphs and the KV caches / layer counters of injected v1 layers."""
        self._graphs = {}
        from ..ops.transformer.inference.ds_transformer import DeepSpeedSelfAttention, DeepSpeedTransformerInference
        DeepSpeedTransformerInference.reset_cache(self.module)
        DeepSpeedTransformerInference.layer_id = 0
        DeepSpeedSelfAttention.num_layers = 0

    def load_model_with_checkpoint(self, r_module):
        """Fill ``r_module`` from the checkpoint given in the config (reference ``engine.py:load_model_with_checkpoint``:
        walks the module tree loading each layer's tensors, slicing them for tensor parallelism)."""
        ckpt = self._config.checkpoint
        self._load_checkpoint(ckpt)
        return r_module

    # ---- BLOOM / MPT helpers: head-sharded alibi and attention bias under tensor parallelism ------------------
    def _head_slice(self):
        if self.mp_world_size <= 1:
            return None
        cfg = getattr(self.module, "config", None)
        heads = next((getattr(cfg, n) for n in ("num_attention_heads", "n_head", "n_heads", "num_heads")
                      if cfg is not None and getattr(cfg, n, None) is not None), None)
        if heads is None:
            return None
        from ..sequence.layer import head_shard_sizes
        sizes = head_shard_sizes(heads, self.mp_world_size)
        r = dist.get_rank(self.mp_group)
        return slice(sum(sizes[:r]), sum(sizes[:r + 1])), heads

    def remove_mask_prepare_for_bloom(self):
        tr = getattr(self.module, "transformer", None)
        if tr is not None and hasattr(tr, "_prepare_attn_mask"):
            tr._prepare_attn_mask = lambda attention_mask, *args, **kwargs: attention_mask

    def _wrap_head_sharded(self, owner, name, head_dim_of):
        """Replace ``owner.<name>`` by a version that returns only this rank's heads of the full-head tensor."""
        hs = self._head_slice()
        fn = getattr(owner, name, None)
        if hs is None or fn is None or getattr(fn, "_ds_head_sharded", False):
            return
        sl, heads = hs

        def sharded(*args, **kwargs):
            out = fn(*args, **kwargs)
            if not torch.is_tensor(out):
                return out
            d = head_dim_of(out, heads)
            if d is None:
                return out
            if out.shape[d] == heads:
                return out.narrow(d, sl.start, sl.stop - sl.start)
            b = out.shape[d] // heads  # [batch * heads, ...] layout (BLOOM)
            shaped = out.reshape(b, heads, *out.shape[d + 1:])
            return shaped[:, sl].reshape(-1, *out.shape[d + 1:])

        sharded._ds_head_sharded = True
        setattr(owner, name, sharded)

    def build_alibi_tensor(self):
        find = lambda out, heads: next((i for i, n in enumerate(out.shape) if n == heads or (i == 0 and n % heads == 0
                                                                                             and n != 1)), None)  # noqa: E731
        for owner_name, fn_name in (("transformer", "build_alibi_tensor"), ("transformer", "build_mpt_alibi_tensor"),
                                    ("model", "get_alibi_mask")):
            owner = getattr(self.module, owner_name, None)
            if owner is not None and hasattr(owner, fn_name):
                self._wrap_head_sharded(owner, fn_name, find)

    def build_attn_bias(self):
        owner = getattr(self.module, "transformer", None)
        if owner is not None and hasattr(owner, "_attn_bias"):
            self._wrap_head_sharded(owner, "_attn_bias",
                                    lambda out, heads: next((i for i, n in enumerate(out.shape) if n == heads), None))

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            mod = self.__dict__.get("_modules", {}).get("module")
            if mod is not None and hasattr(mod, name):
                return getattr(mod, name)
            raise
