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


# ---- the reference's autograd regions by name (``module_inject/layers.py:159-296``) ---------------------------
from .tp import _CopyToTP, _GatherFromTP, _ReduceFromTP, _rk, _shard_sizes, _ws, shard_tensor  # noqa: E402
from .. import comm as dist  # noqa: E402


class RowParallel(torch.autograd.Function):
    """``y = x @ w.T`` on an input-sharded weight, partial sums all-reduced; bias added once after the reduction."""

    @staticmethod
    def forward(ctx, group, input, weight, bias=None):
        ctx.save_for_backward(input, weight)
        ctx.has_bias = bias is not None
        out = torch.matmul(input, weight.t())
        if group is not None and _ws(group) > 1:
            dist.all_reduce(out, group=group)
        return out + bias if bias is not None else out

    @staticmethod
    def backward(ctx, g):
        input, weight = ctx.saved_tensors
        gi = g.matmul(weight)
        g2, i2 = g.reshape(-1, g.shape[-1]), input.reshape(-1, input.shape[-1])
        gb = g2.sum(0) if ctx.has_bias else None
        return None, gi, g2.t().matmul(i2), gb


class ColumnParallel(torch.autograd.Function):
    """``y = x @ w.T + b`` on an output-sharded weight; the input gradient is all-reduced over the group."""

    @staticmethod
    def forward(ctx, group, input, weight, bias=None):
        ctx.save_for_backward(input, weight)
        ctx.group, ctx.has_bias = group, bias is not None
        out = torch.matmul(input, weight.t())
        return out + bias if bias is not None else out

    @staticmethod
    def backward(ctx, g):
        input, weight = ctx.saved_tensors
        gi = g.matmul(weight)
        if ctx.group is not None and _ws(ctx.group) > 1:
            dist.all_reduce(gi, group=ctx.group)
        g2, i2 = g.reshape(-1, g.shape[-1]), input.reshape(-1, input.shape[-1])
        return None, gi, g2.t().matmul(i2), (g2.sum(0) if ctx.has_bias else None)


class AsyncColumnParallel(torch.autograd.Function):
    """:class:`ColumnParallel` whose input-gradient all-reduce runs while the weight-gradient GEMM is computed."""

    @staticmethod
    def forward(ctx, group, input, weight, bias=None):
        ctx.save_for_backward(input, weight)
        ctx.group, ctx.has_bias = group, bias is not None
        out = torch.matmul(input, weight.t())
        return out + bias if bias is not None else out

    @staticmethod
    def backward(ctx, g):
        input, weight = ctx.saved_tensors
        gi = g.matmul(weight).contiguous()
        work = dist.all_reduce(gi, group=ctx.group, async_op=True) \
            if ctx.group is not None and _ws(ctx.group) > 1 else None
        g2, i2 = g.reshape(-1, g.shape[-1]), input.reshape(-1, input.shape[-1])
        gw = g2.t().matmul(i2)
        gb = g2.sum(0) if ctx.has_bias else None
        if work is not None:
            work.wait()
        return None, gi, gw, gb


class GatherFromTensorParallelRegion(torch.autograd.Function):
    """All-gather along the last dim (uneven shards allowed); backward keeps this rank's slice."""

    @staticmethod
    def forward(ctx, input, group, shard_sizes=None):
        world = _ws(group)
        ctx.rank, ctx.world = _rk(group), world
        if world == 1:
            ctx.sizes = [input.shape[-1]]
            return input
        sizes = list(shard_sizes) if shard_sizes is not None else [input.shape[-1]] * world
        ctx.sizes = sizes
        width = max(sizes)
        padded = input if input.shape[-1] == width else F.pad(input, (0, width - input.shape[-1]))
        parts = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(parts, padded.contiguous(), group=group)
        return torch.cat([p[..., :s] for p, s in zip(parts, sizes)], dim=-1)

    @staticmethod
    def backward(ctx, g):
        if ctx.world == 1:
            return g, None, None
        return g.split(ctx.sizes, dim=-1)[ctx.rank].contiguous(), None, None


# ---- layer base + gather context ------------------------------------------------------------------------
class TensorParallel_Layer:
    """Mixin naming what the reference's base class offers on every TP layer: ``is_training_mode``,
    ``gather_params`` / ``_tp_partition`` on the layer's own parameters, and the keep-on-host switch."""
    keep_module_on_host = False

    @classmethod
    def set_keep_module_on_host(cls, value):
        cls.keep_module_on_host = bool(value)

    def is_training_mode(self):
        return is_autotp_training_mode()

    def _tp_dim(self, p):
        return getattr(p, "ds_tp_dim", None) if getattr(p, "ds_tp_sharded", False) else None

    @torch.no_grad()
    def gather_params(self, params_list=None):
        """Replace every sharded parameter's data by the full tensor (all-gather over the TP group)."""
        group = self.mp_group
        for p in (params_list if params_list is not None else list(self.parameters(recurse=False))):
            dim = self._tp_dim(p) if p is not None else None
            if dim is None or _ws(group) == 1:
                continue
            sizes = getattr(p, "ds_tp_sizes", None) or [p.shape[dim]] * _ws(group)
            moved = p.data.movedim(dim, -1)
            full = GatherFromTensorParallelRegion.apply(moved, group, sizes).movedim(-1, dim).contiguous()
            p.ds_tp_shard_shape = tuple(p.shape)
            p.data = full

    @torch.no_grad()
    def _tp_partition(self, params_list=None):
        """Inverse of :meth:`gather_params`: keep this rank's slice again."""
        group = self.mp_group
        for p in (params_list if params_list is not None else list(self.parameters(recurse=False))):
            dim = self._tp_dim(p) if p is not None else None
            if dim is None or _ws(group) == 1 or not hasattr(p, "ds_tp_shard_shape"):
                continue
            sizes = getattr(p, "ds_tp_sizes", None) or [p.ds_tp_shard_shape[dim]] * _ws(group)
            p.data = p.data.split(sizes, dim=dim)[_rk(group)].contiguous()
            del p.ds_tp_shard_shape


class GatherReplacedLayerParams:
    """``with GatherReplacedLayerParams(params_or_module, module):`` - full TP parameters inside the block (for
    export or initialisation), re-partitioned on exit. A no-op when the module is not a TP layer."""

    def __init__(self, params, module, enabled=True):
        self.enabled = enabled and hasattr(module, "gather_params")
        self.module = module
        if isinstance(params, torch.nn.Module):
            params = list(params.parameters(recurse=False))
        elif torch.is_tensor(params):
            params = [params]
        self.params = list(params) if params is not None else None

    def __enter__(self):
        if self.enabled:
            self.module.gather_params(self.params)
        return self

    def __exit__(self, *exc):
        if self.enabled:
            self.module._tp_partition(self.params)
        return False


for _cls in (LinearLayer, LinearAllreduce):  # the mixin's methods on the layers defined in tp.py
    for _name in ("is_training_mode", "_tp_dim", "gather_params", "_tp_partition"):
        if not hasattr(_cls, _name):
            setattr(_cls, _name, getattr(TensorParallel_Layer, _name))


class FusedModuleWrapper:
    """Marks a ``nn.Linear`` holding fused sub-matrices together with their sizes (``(q, k, v)`` or ``(gate, up)``)
    so that :class:`SubParamColumnParallel` shards every sub-matrix on its own."""

    def __init__(self, module, sub_params):
        self.module, self.sub_params = module, list(sub_params)


class SubParamColumnParallel(SubParamLinearLayer):
    """Column-parallel linear over fused sub-parameters; accepts a :class:`FusedModuleWrapper`."""

    def __init__(self, module, mp_group=None, **kw):
        if isinstance(module, FusedModuleWrapper):
            kw.setdefault("sub_params", module.sub_params)
            module = module.module
        spec = getattr(self, "_subparam_layout_spec", None)
        if spec is not None:
            kw.setdefault("sub_params", list(spec[0]))
        super().__init__(module, mp_group, **kw)


class fused_LinearLayer(SubParamColumnParallel):
    """Fused QKV projection: three equal sub-matrices unless ``sub_params`` says otherwise (GQA)."""

    def __init__(self, module, mp_group=None, **kw):
        if not isinstance(module, FusedModuleWrapper) and "sub_params" not in kw and "shape" not in kw:
            kw["sub_params"] = [module.weight.shape[0] // 3] * 3
        super().__init__(module, mp_group, **kw)


class GateUpPack_LinearLayer(SubParamColumnParallel):
    """ChatGLM-style packed ``gate|up`` projection: the two halves are sharded separately."""

    def __init__(self, module, mp_group=None, **kw):
        half = module.weight.shape[0] // 2
        self._subparam_layout_spec = ((half, half), "mlp")
        super().__init__(module, mp_group, **kw)


class _Conv1DAsLinear:
    """View of a GPT-2 ``Conv1D`` (weight ``[in, out]``) with the ``nn.Linear`` orientation."""

    def __init__(self, conv):
        self.weight = nn.Parameter(conv.weight.data.t().contiguous(), requires_grad=conv.weight.requires_grad)
        self.bias = conv.bias


class conv_LinearLayer(LinearLayer):
    """Column-parallel layer built from a ``Conv1D`` (``x @ w + b`` with ``w [in, out]``)."""

    def __init__(self, module, mp_group=None, **kw):
        super().__init__(_Conv1DAsLinear(module), mp_group, **kw)


class Conv_LinearALlreduce(LinearAllreduce):
    """Row-parallel layer built from a ``Conv1D``."""

    def __init__(self, module, mp_group=None, **kw):
        super().__init__(_Conv1DAsLinear(module), mp_group, **kw)


def shard_value_with_share_qk(weight, bias, rank, world_size, shard_value=True, num_kv_heads=None):
    """Yuan-2 localized-filtering attention shares one projection for Q and K: neighbouring Q head pairs map to one V
    head, and K heads sit ``num_heads/2`` further. Returns this rank's V rows (``shard_value``) or output-projection
    columns (reference ``module_inject/fusedqkv_utils.py:214``)."""
    dim = 0 if shard_value else 1
    total = weight.shape[dim]
    heads = num_kv_heads if num_kv_heads is not None else _YUAN["num_kv_heads"]
    if heads is None:
        raise RuntimeError("set_num_kv_heads() must be called before sharding Yuan layers")
    assert heads % world_size == 0, f"{heads} heads do not divide over {world_size} ranks"
    if world_size > heads // 2:
        raise RuntimeError(f"world_size {world_size} is larger than half of num_heads {heads}")
    hd = total // heads
    per = heads // world_size
    ids = [(rank * per + 2 * i) // 2 for i in range((per + 1) // 2)]
    ids += [i + heads // 2 for i in ids]
    w = torch.cat([weight.narrow(dim, h * hd, hd) for h in ids], dim=dim).contiguous()
    if bias is None:
        return w, None
    if shard_value:
        return w, torch.cat([bias.data[h * hd:(h + 1) * hd] for h in ids], dim=0).contiguous()
    return w, bias.data / float(world_size)


_YUAN = {"num_kv_heads": None}


def set_num_kv_heads(n):
    _YUAN["num_kv_heads"] = n


def get_num_kv_heads():
    return _YUAN["num_kv_heads"]


class Yuan_LinearLayer(LinearLayer):
    """Value projection of Yuan-2 with the shared-QK head mapping."""

    def __init__(self, module, mp_group=None, num_kv_heads=None, **kw):
        w, b = shard_value_with_share_qk(module.weight.data, module.bias, _rk(mp_group), _ws(mp_group), True,
                                         num_kv_heads)
        nn.Module.__init__(self)
        self.mp_group, self.gather_output, self.mode = mp_group, False, "allreduce"
        self.tp_world, self.tp_rank, self.sub_params = _ws(mp_group), _rk(mp_group), None
        self.weight = nn.Parameter(w, requires_grad=module.weight.requires_grad)
        self.bias = nn.Parameter(b, requires_grad=module.bias.requires_grad) if b is not None else None
        for p in self.parameters():
            p.tensor_model_parallel, p.ds_tp_sharded, p.ds_tp_dim = True, True, 0
        self.in_features, self.out_features = self.weight.shape[1], self.weight.shape[0]
        self.full_out_features = module.weight.shape[0]

    def gather_params(self, params_list=None):
        raise RuntimeError("Yuan shared-QK tensor parallelism selects non-contiguous head groups; gathering the full "
                           "parameter is not supported")


class Yuan_LinearAllreduce(LinearAllreduce):
    """Output projection of Yuan-2 (columns picked with the same head mapping; bias divided over the ranks is NOT
    needed here: it is added once after the reduction)."""

    def __init__(self, module, mp_group=None, num_kv_heads=None, **kw):
        w, _ = shard_value_with_share_qk(module.weight.data, None, _rk(mp_group), _ws(mp_group), False, num_kv_heads)
        nn.Module.__init__(self)
        self.mp_group, self.mode = mp_group, "allreduce"
        self.tp_world, self.tp_rank = _ws(mp_group), _rk(mp_group)
        self.weight = nn.Parameter(w, requires_grad=module.weight.requires_grad)
        self.weight.tensor_model_parallel, self.weight.ds_tp_sharded, self.weight.ds_tp_dim = True, True, 1
        self.bias = nn.Parameter(module.bias.data.clone(), requires_grad=module.bias.requires_grad) \
            if module.bias is not None else None
        self.in_features, self.out_features = self.weight.shape[1], self.weight.shape[0]

    def gather_params(self, params_list=None):
        raise RuntimeError("Yuan shared-QK tensor parallelism selects non-contiguous head groups; gathering the full "
                           "parameter is not supported")


# ---- convolutions -----------------------------------------------------------------------------------------
class TensorParallelConv2d(nn.Module):
    """``nn.Conv2d`` sharded over output channels (``shard_by_oc``, every conv but the last of a block) or input
    channels (the last one, followed by an all-reduce in :class:`TensorParallelIcShardConv2d`). Channels that do not
    divide evenly go to the last ranks, like the reference (``module_inject/layers.py:1267``)."""

    def __init__(self, conv, rank, world_size, shard_by_oc, group=None):
        super().__init__()
        self.rank, self.world_size, self.shard_by_oc, self.group = rank, world_size, shard_by_oc, group
        total = conv.weight.shape[0] if shard_by_oc else conv.weight.shape[1]
        bounds = [0]
        for i in range(world_size - 1, -1, -1):
            bounds.append(bounds[-1] + total // world_size + (1 if i < total % world_size else 0))
        lo, hi = bounds[rank], bounds[rank + 1]
        if shard_by_oc:
            w = conv.weight.data[lo:hi]
            b = conv.bias.data[lo:hi] if conv.bias is not None else None
        else:
            assert conv.groups == 1, "input-channel sharding of a grouped convolution is not supported"
            w = conv.weight.data[:, lo:hi]
            b = conv.bias.data / float(world_size) if conv.bias is not None else None
        self.conv = nn.Conv2d(w.shape[1] * conv.groups, w.shape[0], conv.kernel_size, conv.stride, conv.padding,
                              conv.dilation, conv.groups, b is not None, conv.padding_mode, device="meta")
        self.conv.weight = nn.Parameter(w.clone().contiguous(), requires_grad=conv.weight.requires_grad)
        if b is not None:
            self.conv.bias = nn.Parameter(b.clone(), requires_grad=conv.bias.requires_grad)
        for p in self.conv.parameters():
            p.tensor_model_parallel = True
        self.channel_range = (lo, hi)

    def forward(self, input):
        return self.conv(input)


class TensorParallelOcShardConv2d(TensorParallelConv2d):

    def __init__(self, conv, rank, world_size, group=None):
        super().__init__(conv, rank, world_size, True, group)

    def forward(self, input):
        if self.world_size > 1 and torch.is_grad_enabled() and input.requires_grad:
            input = _CopyToTP.apply(input, self.group)  # replicated input: its gradient is a sum over the ranks
        return self.conv(input)


class TensorParallelIcShardConv2d(TensorParallelConv2d):

    def __init__(self, conv, rank, world_size, group=None):
        super().__init__(conv, rank, world_size, False, group)

    def forward(self, input):
        out = self.conv(input)
        if self.world_size > 1:
            out = _ReduceFromTP.apply(out, self.group)
        return out


# ---- norms / embeddings ------------------------------------------------------------------------------------
class RMSNormalize(nn.Module):
    """RMSNorm holder for checkpoint loading; on a GPU the fused kernel (``ops/fused.py``) does the work."""

    def __init__(self, dim=None, dtype=torch.float, eps=1e-5, weight=None):
        super().__init__()
        self.weight = weight if weight is not None else nn.Parameter(torch.ones(dim, dtype=dtype))
        self.eps = eps

    def forward(self, hidden_states):
        if hidden_states.is_cuda and hidden_states.dtype == self.weight.dtype and \
                hidden_states.dtype in (torch.bfloat16, torch.float16):
            from ..ops import fused
            return fused.rms_norm(hidden_states, self.weight, self.eps)
        var = hidden_states.to(torch.float32).pow(2).mean(-1, keepdim=True)
        h = hidden_states * torch.rsqrt(var + self.eps)
        if self.weight.dtype in (torch.float16, torch.bfloat16):
            h = h.to(self.weight.dtype)
        return h * self.weight


class OPTEmbedding(EmbeddingLayer):
    """OPT's learned positional embedding: positions come from the attention mask and are offset by 2."""

    def __init__(self, weight_shape=None, weight=None, bias=None):
        self.offset = 2
        super().__init__(weight_shape, weight=weight)

    def forward(self, attention_mask, past_key_values_length=0, position_ids=0):
        attention_mask = attention_mask.long()
        positions = (torch.cumsum(attention_mask, dim=1).type_as(attention_mask) * attention_mask).long() - 1
        return super().forward(positions[:, past_key_values_length:] + self.offset)
