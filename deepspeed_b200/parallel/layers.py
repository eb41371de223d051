"""Remaining layer names of the reference ``module_inject/layers.py``: embedding / norm wrappers used when weights
are loaded from sharded checkpoints, sub-parameter (fused QKV / gate-up) linears, the AutoTP mode switch."""
from enum import Enum

import torch
import torch.nn as nn
import torch.nn.functional as F

from .tp import LinearAllreduce, LinearLayer


class AUTOTP_MODE(Enum):
    TRAINING = "TRAINING"
    INFERENCE = "INFERENCE"


DEEPSPEED_AUTOTP_MODE = AUTOTP_MODE.INFERENCE


def is_autotp_training_mode():
    return DEEPSPEED_AUTOTP_MODE == AUTOTP_MODE.TRAINING


def set_autotp_mode(training=False):
    """Training mode keeps autograd-aware collectives and trainable shards; inference mode freezes them."""
    global DEEPSPEED_AUTOTP_MODE
    DEEPSPEED_AUTOTP_MODE = AUTOTP_MODE.TRAINING if training else AUTOTP_MODE.INFERENCE


class EmbeddingLayer(nn.Module):
    """Embedding whose table is allocated empty (filled by the checkpoint loader)."""

    def __init__(self, weight_shape=None, dtype=torch.half, weight=None, bias=None):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(weight_shape[0], weight_shape[1], dtype=dtype)) if weight is None \
            else weight
        self.weight.ds_tensor = None

    def forward(self, input):
        return F.embedding(input, self.weight)


class Normalize(nn.Module):
    """LayerNorm holder for checkpoint loading: wraps a given norm or owns fresh weight / bias of ``dim``."""

    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None, bias=None):
        super().__init__()
        if weight is not None:
            self.weight, self.bias = weight, bias
        else:
            norm = nn.LayerNorm(dim, eps=eps).to(dtype)
            self.weight, self.bias = norm.weight, norm.bias
        self.eps = eps

    def forward(self, input):
        return F.layer_norm(input, input.shape[-1:], self.weight, self.bias, eps=self.eps)


class SubParamLinearLayer(LinearLayer):
    """Column-parallel linear over a weight made of fused sub-parameters (``sub_params`` sizes along dim 0: q|k|v,
    gate|up): every rank holds matching slices of each sub-parameter."""

    def __init__(self, module, mp_group, shape=None, partition_dim=0, name=None, **kw):
        sub = kw.pop("sub_params", None)
        if sub is None and shape is not None:
            d = shape[partition_dim]
            sub = list(d) if isinstance(d, (tuple, list)) else [module.weight.shape[0] // d] * d
        super().__init__(module, mp_group, sub_params=sub, **kw)


class SubParamLinearAllreduce(LinearAllreduce):
    """Row-parallel counterpart: the input dim is made of fused sub-parameters."""

    def __init__(self, module, mp_group, shape=None, partition_dim=1, name=None, **kw):
        sub = kw.pop("sub_params", None)
        if sub is None and shape is not None:
            d = shape[partition_dim]
            sub = list(d) if isinstance(d, (tuple, list)) else [module.weight.shape[1] // d] * d
        super().__init__(module, mp_group, sub_params=sub, **kw)
