"""``DeepSpeedHybridEngine``: one engine that trains (ZeRO) and generates (RLHF rollouts) with the same weights.

Reference: ``runtime/hybrid_engine.py:31`` - for generation it swaps transformer blocks for inference containers,
fuses LoRA adapters into the base weights, gathers ZeRO-3 partitions layer by layer and un-does all of it for
training. Here generation runs the training module itself (its kernels are already the fused sm_100a ones), so the
switch consists of: (1) make the full parameters resident (ZeRO-3: one gather of every unit, kept for the whole
``generate`` call instead of per-layer fetch/release per token), (2) fuse LoRA, (3) ``no_grad`` + eval mode +
KV-cached decoding when the module offers ``generate``, otherwise a cached-free greedy/top-p loop, (4) undo."""
import time

import torch

from ..utils.logging import log_dist, logger
from .engine import DeepSpeedEngine


def _lora_pairs(module):
    """(layer, base weight, right, left, scaling) for every LoRA-carrying linear: DeepSpeed-Chat layers
    (``lora_right_weight`` / ``lora_left_weight`` / ``lora_scaling``) and ``deepspeed.linear.LoRAOptimizedLinear``."""
    out = []
    for m in module.modules():
        if hasattr(m, "lora_right_weight") and hasattr(m, "lora_left_weight"):
            out.append((m, m.weight, m.lora_right_weight, m.lora_left_weight, getattr(m, "lora_scaling", 1.0), "chat"))
        elif hasattr(m, "lora_weight_1") and hasattr(m, "lora_weight_2") and getattr(m, "_initialized", False) \
                and m.weight.is_floating_point() and getattr(m, "zero_shards", 1) == 1:
            out.append((m, m.weight, m.lora_weight_1.weight, m.lora_weight_2.weight, m.lora_scaling_factor, "opt"))
    return out


class DeepSpeedHybridEngine(DeepSpeedEngine):
    inference_mp_size = 1

    def __init__(self, args, model, **kwargs):
        super().__init__(args, model, **kwargs)
        cfg = self._config.hybrid_engine
        self.Z3_enabled = self.zero_optimization_stage() == 3
        self.max_out_tokens = cfg.max_out_tokens
        self.release_inference_cache = cfg.release_inference_cache
        self.inference_tp_size = cfg.inference_tp_size
        self._total_batch_size = None
        self._gather_latency = 0.0
        self._generate_latency = 0.0
        self._training_latency = 0.0
        self._total_latency = 0.0
        self._iters = 0
        self._training_start_time = None
        self._t_start = None
        self.is_lora_fused = False
        self._gather_ctx = None
        self._orig_generate = getattr(self.module, "generate", None)
        self._decode_graphs = None
        if getattr(cfg, "enable_cuda_graph", False):
            from .hybrid_engine_graph import DecodeGraphCache, validate_cuda_graph_support
            why = validate_cuda_graph_support(cfg, self.zero_optimization_stage())
            if why is None:
                self._decode_graphs = DecodeGraphCache(self.module.forward, self.max_out_tokens)
            else:
                logger.warning(f"hybrid_engine.enable_cuda_graph ignored: {why}")
        log_dist(f"DeepSpeedHybridEngine: ZeRO-3={self.Z3_enabled}, LoRA layers={len(_lora_pairs(self.module))}", [0])

    # ---- LoRA -----------------------------------------------------------------------------------
    @torch.no_grad()
    def fuse_lora_weight(self):
        if self.is_lora_fused:
            return
        for m, w, a, b, scale, kind in _lora_pairs(self.module):
            delta = (a @ b).t() if kind == "chat" else (b @ a)  # chat: W[out,in] += s * (right[in,r] @ left[r,out])^T
            w.data.add_(delta.to(w.dtype), alpha=scale)
        self.is_lora_fused = True

    @torch.no_grad()
    def unfuse_lora_weight(self):
        if not self.is_lora_fused:
            return
        for m, w, a, b, scale, kind in _lora_pairs(self.module):
            delta = (a @ b).t() if kind == "chat" else (b @ a)
            w.data.sub_(delta.to(w.dtype), alpha=scale)
        self.is_lora_fused = False

    def unfuse_lora_weight_non_pinned(self):
        self.unfuse_lora_weight()

    def retake_inference_cache(self):
        """KV-cache workspace is owned by the module's generate(); nothing persistent to re-acquire."""
        return None

    # ---- parameter residency ----------------------------------------------------------------------
    def _enter_generation(self):
        t0 = time.time()
        if self.Z3_enabled and self.optimizer is not None:
            from .zero.init_context import GatheredParameters
            self._gather_ctx = GatheredParameters(list(self.module.parameters()), modifier_rank=None)
            self._gather_ctx.__enter__()
        self.fuse_lora_weight()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self._gather_latency += time.time() - t0

    def _exit_generation(self):
        self.unfuse_lora_weight()
        if self._gather_ctx is not None:
            self._gather_ctx.__exit__(None, None, None)
            self._gather_ctx = None
        if self.release_inference_cache and torch.cuda.is_available():
            torch.cuda.empty_cache()

    # ---- generation -------------------------------------------------------------------------------
    @torch.no_grad()
    def generate(self, *inputs, **kwargs):
        was_training = self.module.training
        self.module.eval()
        if self._total_batch_size is None and inputs:
            bsz = inputs[0].shape[0] if torch.is_tensor(inputs[0]) else 1
            self._total_batch_size = bsz * self.dp_world_size
        self._t_start = time.time()
        self._enter_generation()
        try:
            t0 = time.time()
            graphs = self._decode_graphs
            if graphs is not None:
                from .hybrid_engine_graph import decode_steps_from_generate_kwargs
                graphs.begin_sequence(decode_steps_from_generate_kwargs(kwargs))
                self.module.forward = graphs  # the decode forwards of this call go through the graph cache
            try:
                if self._orig_generate is not None:
                    out = self._orig_generate(*inputs, **kwargs)
                else:
                    out = self._greedy_generate(*inputs, **kwargs)
            finally:
                if graphs is not None:
                    del self.module.forward  # back to the class's forward
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            self._generate_latency += time.time() - t0
        finally:
            self._exit_generation()
            if was_training:
                self.module.train()
        return out

    def _greedy_generate(self, input_ids, max_new_tokens=None, max_length=None, temperature=1.0, top_p=1.0,
                         do_sample=False, eos_token_id=None, pad_token_id=0, attention_mask=None, **_):
        """Fallback for modules without ``generate``: recompute-all decoding (no cache), greedy or nucleus."""
        if max_new_tokens is None:
            max_new_tokens = (max_length - input_ids.shape[1]) if max_length else self.max_out_tokens
        ids = input_ids
        finished = torch.zeros(ids.shape[0], dtype=torch.bool, device=ids.device)
        for _ in range(max_new_tokens):
            out = self.module(ids)
            logits = out.logits if hasattr(out, "logits") else (out[0] if isinstance(out, (tuple, list)) else out)
            logits = logits[:, -1, :].float()
            if do_sample:
                nxt = sample_top_p(logits, temperature, top_p)
            else:
                nxt = logits.argmax(-1)
            nxt = torch.where(finished, torch.full_like(nxt, pad_token_id), nxt)
            ids = torch.cat([ids, nxt[:, None]], dim=1)
            if eos_token_id is not None:
                finished |= nxt == eos_token_id
                if bool(finished.all()):
                    break
        return ids

    # ---- mode switches ----------------------------------------------------------------------------
    def eval(self):
        if self._t_start is not None and self._training_start_time is not None:
            self._training_latency += time.time() - self._training_start_time
            self._total_latency += time.time() - self._t_start
            self._iters += 1
            if self._iters % max(1, self.steps_per_print()) == 0:
                log_dist(f"hybrid engine: gather {self._gather_latency:.2f}s generate {self._generate_latency:.2f}s "
                         f"train {self._training_latency:.2f}s over {self._iters} iterations", [0])
        self._training_start_time = None
        return super().eval()

    def train(self, mode=True):
        if mode:
            self._training_start_time = time.time()
        return super().train(mode)

    def step(self, lr_kwargs=None):
        super().step(lr_kwargs=lr_kwargs)


def sample_top_p(logits, temperature=1.0, top_p=1.0):
    """Nucleus sampling: keep the smallest set of tokens whose probability mass reaches ``top_p``."""
    if temperature != 1.0:
        logits = logits / max(temperature, 1e-5)
    probs = torch.softmax(logits, dim=-1)
    if top_p < 1.0:
        sp, si = probs.sort(dim=-1, descending=True)
        keep = (sp.cumsum(-1) - sp) < top_p
        sp = sp * keep
        sp = sp / sp.sum(-1, keepdim=True)
        return si.gather(-1, torch.multinomial(sp, 1)).squeeze(-1)
    return torch.multinomial(probs, 1).squeeze(-1)
