"""Mixtral-style sparse MoE decoder (config 5 of BASELINE.json: Mixtral 8x7B, expert parallel all-to-all).

Attention / norms / loss are shared with :mod:`.llama`; every layer's FFN is an :class:`AutoEPMoELayer` (dropless
top-2 routing, grouped-GEMM SwiGLU experts, expert-parallel dispatch/combine)."""
from dataclasses import dataclass

import torch
import torch.nn as nn

from ..parallel.auto_ep import AutoEPMoELayer
from .llama import LlamaAttention, LlamaConfig, LMHead, RMSNorm
from ..ops import fused


@dataclass
class MixtralConfig(LlamaConfig):
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 14336
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 8
    max_position_embeddings: int = 32768
    rope_theta: float = 1e6
    num_local_experts: int = 8
    num_experts_per_tok: int = 2
    ep_size: int = 1
    router_aux_loss_coef: float = 0.02

    @staticmethod
    def tiny(**kw):
        base = dict(vocab_size=512, hidden_size=128, intermediate_size=256, num_hidden_layers=2,
                    num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=256, num_local_experts=4,
                    num_experts_per_tok=2)
        base.update(kw)
        return MixtralConfig(**base)


class MixtralDecoderLayer(nn.Module):

    def __init__(self, cfg: MixtralConfig, device=None):
        super().__init__()
        self.input_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self.self_attn = LlamaAttention(cfg, device)
        self.post_attention_layernorm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self.block_sparse_moe = AutoEPMoELayer(cfg.hidden_size, cfg.intermediate_size, cfg.num_local_experts,
                                               cfg.num_experts_per_tok, cfg.ep_size, cfg.dtype, device,
                                               aux_loss_coef=cfg.router_aux_loss_coef)

    def forward(self, h, cos, sin, position_ids=None, sp_attention=None):
        a = self.self_attn(self.input_layernorm(h), cos, sin, position_ids, sp_attention)
        x, h = self.post_attention_layernorm(a, residual=h)
        return h + self.block_sparse_moe(x)


class MixtralForCausalLM(nn.Module):

    def __init__(self, cfg: MixtralConfig, device=None):
        super().__init__()
        self.cfg = cfg
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size, dtype=cfg.dtype, device=device)
        self.layers = nn.ModuleList([MixtralDecoderLayer(cfg, device) for _ in range(cfg.num_hidden_layers)])
        self.norm = RMSNorm(cfg.hidden_size, cfg.rms_norm_eps, cfg.dtype, device)
        self.lm_head = LMHead(cfg, device)
        self._rope = None
        self.sp_attention = None
        with torch.no_grad():
            for n, p in self.named_parameters():
                if p.dim() >= 2:
                    p.normal_(0.0, cfg.init_std)
                elif "norm" in n:
                    p.fill_(1.0)

    def set_deepspeed_parallelism(self, use_data_before_expert_parallel_=False):
        for l in self.layers:
            l.block_sparse_moe.set_deepspeed_parallelism(use_data_before_expert_parallel_)

    def forward(self, input_ids, labels=None, position_ids=None):
        B, S = input_ids.shape
        if self._rope is None or self._rope[0].shape[0] < S or self._rope[0].device != input_ids.device:
            self._rope = fused.rope_tables(max(S, 1), self.cfg.head_dim, self.cfg.rope_theta, input_ids.device)
        cos, sin = self._rope
        h = self.embed_tokens(input_ids)
        ckpt = self.cfg.gradient_checkpointing and self.training and torch.is_grad_enabled()
        for layer in self.layers:
            if ckpt:
                h = torch.utils.checkpoint.checkpoint(layer, h, cos, sin, position_ids, self.sp_attention,
                                                      use_reentrant=False)
            else:
                h = layer(h, cos, sin, position_ids, self.sp_attention)
        h = self.norm(h)
        if labels is None:
            return self.lm_head(h)
        tgt = torch.full_like(labels, -100)
        tgt[:, :-1] = labels[:, 1:]
        loss = self.lm_head(h, tgt)
        aux = [l.block_sparse_moe.l_aux for l in self.layers if l.block_sparse_moe.l_aux is not None]
        return loss + sum(aux) if aux else loss
