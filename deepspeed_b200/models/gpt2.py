"""GPT-2 (config 1 of BASELINE.json: GPT-2 small, ZeRO-1, CPU/gloo plumbing run; also a GPU model)."""
from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import fused


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    layer_norm_epsilon: float = 1e-5
    dropout: float = 0.0
    gradient_checkpointing: bool = False
    dtype: torch.dtype = torch.float32

    @staticmethod
    def small(**kw):
        return GPT2Config(**kw)

    @staticmethod
    def tiny(**kw):
        base = dict(vocab_size=256, n_positions=64, n_embd=64, n_layer=2, n_head=4)
        base.update(kw)
        return GPT2Config(**base)


class LayerNorm(nn.Module):

    def __init__(self, dim, eps, dtype):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim, dtype=dtype))
        self.bias = nn.Parameter(torch.zeros(dim, dtype=dtype))
        self.eps = eps

    def forward(self, x):
        return fused.layer_norm(x, self.weight, self.bias, self.eps)


class GPT2Block(nn.Module):

    def __init__(self, cfg):
        super().__init__()
        d = cfg.n_embd
        self.ln_1 = LayerNorm(d, cfg.layer_norm_epsilon, cfg.dtype)
        self.c_attn = nn.Linear(d, 3 * d, dtype=cfg.dtype)
        self.c_proj = nn.Linear(d, d, dtype=cfg.dtype)
        self.ln_2 = LayerNorm(d, cfg.layer_norm_epsilon, cfg.dtype)
        self.c_fc = nn.Linear(d, 4 * d, dtype=cfg.dtype)
        self.c_proj2 = nn.Linear(4 * d, d, dtype=cfg.dtype)
        self.n_head = cfg.n_head
        self.drop = cfg.dropout

    def forward(self, x):
        B, S, D = x.shape
        q, k, v = self.c_attn(self.ln_1(x)).split(D, dim=-1)
        sh = lambda t: t.view(B, S, self.n_head, D // self.n_head).transpose(1, 2)  # noqa: E731
        a = F.scaled_dot_product_attention(sh(q), sh(k), sh(v), is_causal=True,
                                           dropout_p=self.drop if self.training else 0.0)
        x = x + self.c_proj(a.transpose(1, 2).reshape(B, S, D))
        return x + self.c_proj2(F.gelu(self.c_fc(self.ln_2(x)), approximate="tanh"))


class GPT2LMHeadModel(nn.Module):

    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd, dtype=cfg.dtype)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd, dtype=cfg.dtype)
        self.h = nn.ModuleList([GPT2Block(cfg) for _ in range(cfg.n_layer)])
        self.ln_f = LayerNorm(cfg.n_embd, cfg.layer_norm_epsilon, cfg.dtype)
        self.apply(self._init)

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward(self, input_ids, labels=None):
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device)
        x = self.wte(input_ids) + self.wpe(pos)[None]
        for blk in self.h:
            if self.cfg.gradient_checkpointing and self.training:
                x = torch.utils.checkpoint.checkpoint(blk, x, use_reentrant=False)
            else:
                x = blk(x)
        x = self.ln_f(x)
        if labels is None:
            return F.linear(x, self.wte.weight)
        tgt = torch.full_like(labels, -100)
        tgt[:, :-1] = labels[:, 1:]
        return fused.linear_cross_entropy(x, self.wte.weight, tgt)
