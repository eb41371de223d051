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

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            mod = self.__dict__.get("_modules", {}).get("module")
            if mod is not None and hasattr(mod, name):
                return getattr(mod, name)
            raise
