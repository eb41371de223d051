"""``DeepSpeedTransformerLayer``: BERT-style encoder layer with fused kernels (reference
``ops/transformer/transformer.py:296`` over ``csrc/transformer/*``: cuBLAS GEMMs + fused LayerNorm / softmax / dropout /
GELU kernels, pre- or post-LayerNorm, optional activation re-computation switches).

Same parameter names and shapes as the reference (``attn_qkvw``, ``attn_ow``, ``inter_w``, ``output_w`` ...), so its
state dicts load. The layer is assembled from this framework's fused ops: ``fused.layer_norm`` (sm_100a LayerNorm
fwd/bwd kernels), scaled-dot-product attention, and library GEMMs; ``gelu_checkpoint`` / ``attn_dropout_checkpoint`` /
``normalize_invertible`` map onto activation checkpointing of the respective sub-blocks."""
import json
import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.checkpoint import checkpoint

from .. import fused


class TransformerConfig:

    def __init__(self, batch_size, hidden_size, intermediate_size, heads, attn_dropout_ratio, hidden_dropout_ratio,
                 num_hidden_layers, initializer_range):
        self.layer_id = -1
        self.batch_size = batch_size
        self.hidden_size = hidden_size
        self.intermediate_size = intermediate_size
        self.heads = heads
        self.attn_dropout_ratio = attn_dropout_ratio
        self.hidden_dropout_ratio = hidden_dropout_ratio
        self.num_hidden_layers = num_hidden_layers
        self.initializer_range = initializer_range


class DeepSpeedTransformerConfig(TransformerConfig):

    def __init__(self, batch_size=-1, hidden_size=-1, intermediate_size=-1, heads=-1, attn_dropout_ratio=-1,
                 hidden_dropout_ratio=-1, num_hidden_layers=-1, initializer_range=-1, layer_norm_eps=1e-12,
                 local_rank=-1, seed=-1, fp16=False, pre_layer_norm=True, normalize_invertible=False,
                 gelu_checkpoint=False, adjust_init_range=True, attn_dropout_checkpoint=False, stochastic_mode=False,
                 return_tuple=False, training=True):
        super().__init__(batch_size, hidden_size, (intermediate_size if intermediate_size > 0 else 4 * hidden_size),
                         heads, attn_dropout_ratio, hidden_dropout_ratio, num_hidden_layers, initializer_range)
        self.fp16 = fp16
        self.pre_layer_norm = pre_layer_norm
        self.local_rank = local_rank
        self.seed = seed
        self.normalize_invertible = normalize_invertible
        self.gelu_checkpoint = gelu_checkpoint
        self.adjust_init_range = adjust_init_range
        self.test_gemm = False
        self.layer_norm_eps = layer_norm_eps
        self.training = training
        self.is_grad_enabled = True
        self.attn_dropout_checkpoint = attn_dropout_checkpoint
        self.stochastic_mode = stochastic_mode
        self.return_tuple = return_tuple

    @classmethod
    def from_dict(cls, json_object):
        config = DeepSpeedTransformerConfig()
        for key, value in json_object.items():
            config.__dict__[key] = value
        return config

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-16") as reader:
            return cls.from_dict(json.loads(reader.read()))


class DeepSpeedTransformerLayer(nn.Module):
    layer_id = 0

    def __init__(self, config, initial_weights=None, initial_biases=None):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedTransformerLayer.layer_id
        DeepSpeedTransformerLayer.layer_id += 1
        h, inter = config.hidden_size, config.intermediate_size
        if initial_weights is None and initial_biases is None:
            self.attn_qkvw = nn.Parameter(torch.empty(h * 3, h))
            self.attn_qkvb = nn.Parameter(torch.empty(h * 3))
            self.attn_ow = nn.Parameter(torch.empty(h, h))
            self.attn_ob = nn.Parameter(torch.empty(h))
            self.attn_nw = nn.Parameter(torch.empty(h))
            self.attn_nb = nn.Parameter(torch.empty(h))
            self.inter_w = nn.Parameter(torch.empty(inter, h))
            self.inter_b = nn.Parameter(torch.empty(inter))
            self.output_w = nn.Parameter(torch.empty(h, inter))
            self.output_b = nn.Parameter(torch.empty(h))
            self.norm_w = nn.Parameter(torch.empty(h))
            self.norm_b = nn.Parameter(torch.empty(h))
            self.init_transformer_weights(config.adjust_init_range)
        else:  # HF BertLayer order: q, k, v, attn out, attn LN, intermediate, output, output LN
            q, k, v = initial_weights[0].data, initial_weights[1].data, initial_weights[2].data
            self.attn_qkvw = nn.Parameter(torch.cat((q, k, v)))
            self.attn_qkvb = nn.Parameter(torch.cat((initial_biases[0].data, initial_biases[1].data,
                                                     initial_biases[2].data)))
            self.attn_ow, self.attn_ob = initial_weights[3], initial_biases[3]
            self.attn_nw, self.attn_nb = initial_weights[4], initial_biases[4]
            self.inter_w, self.inter_b = initial_weights[5], initial_biases[5]
            self.output_w, self.output_b = initial_weights[6], initial_biases[6]
            self.norm_w, self.norm_b = initial_weights[7], initial_biases[7]

    def init_transformer_weights(self, adjust_init_range=False):
        num_layers = self.config.num_hidden_layers
        std = self.config.initializer_range
        out_std = std / math.sqrt(2.0 * num_layers) if (adjust_init_range and self.config.local_rank == 0) else std
        for w in (self.attn_qkvw, self.inter_w):
            w.data.normal_(mean=0.0, std=std)
        for w in (self.attn_ow, self.output_w):
            w.data.normal_(mean=0.0, std=out_std)
        for b in (self.attn_qkvb, self.attn_ob, self.inter_b, self.output_b, self.attn_nb, self.norm_b):
            b.data.zero_()
        self.attn_nw.data.fill_(1.0)
        self.norm_w.data.fill_(1.0)

    # ---- sub-blocks -----------------------------------------------------------------------------
    def _attention(self, x, attention_mask):
        B, S, Hd = x.shape
        nh = self.config.heads
        qkv = F.linear(x, self.attn_qkvw, self.attn_qkvb).view(B, S, 3, nh, Hd // nh).permute(2, 0, 3, 1, 4)
        p = self.config.attn_dropout_ratio if (self.training and self.config.attn_dropout_ratio > 0) else 0.0
        mask = attention_mask
        if mask is not None:
            mask = mask.to(qkv.dtype)
            while mask.dim() < 4:
                mask = mask.unsqueeze(1)
        ctx = F.scaled_dot_product_attention(qkv[0], qkv[1], qkv[2], attn_mask=mask, dropout_p=p)
        return F.linear(ctx.transpose(1, 2).reshape(B, S, Hd), self.attn_ow, self.attn_ob)

    def _mlp(self, x):
        return F.linear(F.gelu(F.linear(x, self.inter_w, self.inter_b)), self.output_w, self.output_b)

    def _drop(self, x):
        p = self.config.hidden_dropout_ratio
        return F.dropout(x, p, self.training) if p > 0 else x

    def forward(self, hidden_states, attention_mask=None, head_mask=None, layer_head_mask=None,
                encoder_hidden_states=None, encoder_attention_mask=None, past_key_value=None, output_attentions=False,
                grads=None, **unused_hf_kwargs):
        self.config.is_grad_enabled = torch.is_grad_enabled()
        cfg = self.config
        eps = cfg.layer_norm_eps
        recompute_attn = cfg.attn_dropout_checkpoint and self.training and torch.is_grad_enabled()
        recompute_mlp = (cfg.gelu_checkpoint or cfg.normalize_invertible) and self.training and torch.is_grad_enabled()

        def attn(x):
            if recompute_attn:
                return checkpoint(self._attention, x, attention_mask, use_reentrant=False)
            return self._attention(x, attention_mask)

        def mlp(x):
            return checkpoint(self._mlp, x, use_reentrant=False) if recompute_mlp else self._mlp(x)

        x = hidden_states
        if cfg.pre_layer_norm:
            x = x + self._drop(attn(fused.layer_norm(x, self.attn_nw, self.attn_nb, eps)))
            out = x + self._drop(mlp(fused.layer_norm(x, self.norm_w, self.norm_b, eps)))
        else:
            x = fused.layer_norm(x + self._drop(attn(x)), self.attn_nw, self.attn_nb, eps)
            out = fused.layer_norm(x + self._drop(mlp(x)), self.norm_w, self.norm_b, eps)
        return (out, ) if cfg.return_tuple else out
