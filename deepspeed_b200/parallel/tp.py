"""AutoTP: tensor-parallel layers and automatic module replacement for training and inference.

Reference: ``module_inject/layers.py`` (``LinearLayer`` column-parallel :752-930, ``LinearAllreduce`` row-parallel
:159-296 with all-reduce in forward, fused-QKV / GQA / gate-up sub-parameter sharding :942-1264, lm-head / embedding
:1642-1880), ``module_inject/auto_tp.py:198`` (policy: which linears are row-parallel), ``autotp_config.py``,
``tp_plan_converter.py`` (HF ``tp_plan`` import), ``runtime/tensor_parallel/config.py``.

Differences by design: the row-parallel forward can end in a fused GEMM->reduce-scatter and the following
column-parallel layer can start with all-gather->GEMM (sequence-parallel activations between, half the bytes of an
all-reduce and no redundant norm/elementwise work) - ``mode="rs_ag"``; ``mode="allreduce"`` reproduces the reference.
"""
import re
from typing import Dict, List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import comm as dist
from ..utils.logging import log_dist, logger


def _ws(group):
    return dist.get_world_size(group) if (dist.is_initialized() and group is not None) else 1


def _rk(group):
    return dist.get_rank(group) if (dist.is_initialized() and group is not None) else 0


class _CopyToTP(torch.autograd.Function):
    """Identity forward, all-reduce backward (input of a column-parallel layer)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, g):
        if _ws(ctx.group) > 1:
            g = g.contiguous()
            dist.all_reduce(g, group=ctx.group)
        return g, None


class _ReduceFromTP(torch.autograd.Function):
    """All-reduce forward, identity backward (output of a row-parallel layer)."""

    @staticmethod
    def forward(ctx, x, group):
        if _ws(group) > 1:
            x = x.contiguous()
            dist.all_reduce(x, group=group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None


class _GatherFromTP(torch.autograd.Function):
    """All-gather the last dim forward, split backward (``gather_output``)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group, ctx.world, ctx.rank = group, _ws(group), _rk(group)
        if ctx.world == 1:
            return x
        parts = [torch.empty_like(x) for _ in range(ctx.world)]
        dist.all_gather(parts, x.contiguous(), group=group)
        return torch.cat(parts, dim=-1)

    @staticmethod
    def backward(ctx, g):
        if ctx.world == 1:
            return g, None
        return g.chunk(ctx.world, dim=-1)[ctx.rank].contiguous(), None


class _ReduceScatterSeq(torch.autograd.Function):
    """Reduce-scatter along the sequence dim (dim 1) forward, all-gather backward."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group, ctx.world = group, _ws(group)
        if ctx.world == 1:
            return x
        xs = x.transpose(0, 1).contiguous()  # [S, B, H]
        out = torch.empty((xs.shape[0] // ctx.world, ) + tuple(xs.shape[1:]), dtype=x.dtype, device=x.device)
        dist.reduce_scatter_tensor(out, xs, group=group)
        return out.transpose(0, 1).contiguous()

    @staticmethod
    def backward(ctx, g):
        if ctx.world == 1:
            return g, None
        gs = g.transpose(0, 1).contiguous()
        out = torch.empty((gs.shape[0] * ctx.world, ) + tuple(gs.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, gs, group=ctx.group)
        return out.transpose(0, 1).contiguous(), None


class _AllGatherSeq(torch.autograd.Function):
    """All-gather along the sequence dim forward, reduce-scatter backward."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group, ctx.world = group, _ws(group)
        if ctx.world == 1:
            return x
        xs = x.transpose(0, 1).contiguous()
        out = torch.empty((xs.shape[0] * ctx.world, ) + tuple(xs.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, xs, group=ctx.group)
        return out.transpose(0, 1).contiguous()

    @staticmethod
    def backward(ctx, g):
        if ctx.world == 1:
            return g, None
        gs = g.transpose(0, 1).contiguous()
        out = torch.empty((gs.shape[0] // ctx.world, ) + tuple(gs.shape[1:]), dtype=g.dtype, device=g.device)
        dist.reduce_scatter_tensor(out, gs, group=ctx.group)
        return out.transpose(0, 1).contiguous(), None


def _shard_sizes(total, world, grain=1):
    """Possibly uneven shard sizes in multiples of ``grain`` (reference ``get_shard_size_list``)."""
    units = total // grain
    base, extra = divmod(units, world)
    sizes = [(base + (1 if r < extra else 0)) * grain for r in range(world)]
    sizes[-1] += total - sum(sizes)
    return sizes


def shard_tensor(w, dim, group, sub_params: Optional[List[int]] = None, grain=1):
    """This rank's slice of ``w`` along ``dim``. ``sub_params``: sizes of fused sub-matrices (q|k|v, gate|up) that are
    each sharded separately so that every rank holds matching slices of all of them."""
    world, rank = _ws(group), _rk(group)
    if world == 1:
        return w
    if sub_params:
        pieces = []
        for sub in w.split(list(sub_params), dim=dim):
            sizes = _shard_sizes(sub.shape[dim], world, grain)
            pieces.append(sub.split(sizes, dim=dim)[rank])
        return torch.cat(pieces, dim=dim).contiguous()
    sizes = _shard_sizes(w.shape[dim], world, grain)
    return w.split(sizes, dim=dim)[rank].contiguous()


class LinearLayer(nn.Module):
    """Column-parallel linear: weight ``[out/tp, in]``. Input replicated (or sequence-sharded with ``mode='rs_ag'``)."""

    def __init__(self, module: nn.Linear, mp_group, sub_params=None, gather_output=False, mode="allreduce", grain=1,
                 row_index=None):
        super().__init__()
        self.mp_group, self.gather_output, self.mode = mp_group, gather_output, mode
        self.tp_world, self.tp_rank = _ws(mp_group), _rk(mp_group)
        self.sub_params = sub_params
        self.row_index = row_index  # explicit rows of the full weight (fused-QKV layouts, parallel/fusedqkv_utils.py)
        if row_index is not None and self.tp_world > 1:
            idx = row_index.to(module.weight.device)
            w = module.weight.data.index_select(0, idx).contiguous()
            b = module.bias.data.index_select(0, idx).contiguous() if module.bias is not None else None
        else:
            w = shard_tensor(module.weight.data, 0, mp_group, sub_params, grain)
            b = shard_tensor(module.bias.data, 0, mp_group, sub_params, grain) if module.bias is not None else None
        self.weight = nn.Parameter(w, requires_grad=module.weight.requires_grad)
        self.bias = None
        if b is not None:
            self.bias = nn.Parameter(b, requires_grad=module.bias.requires_grad)
        for p in self.parameters():
            p.tensor_model_parallel = True
            p.ds_tp_sharded = True
            p.ds_tp_dim = 0
        self.in_features, self.out_features = self.weight.shape[1], self.weight.shape[0]
        self.full_out_features = module.weight.shape[0]

    def forward(self, x):
        if self.mode == "rs_ag":
            x = _AllGatherSeq.apply(x, self.mp_group)   # fused AG->GEMM kernel target
        else:
            x = _CopyToTP.apply(x, self.mp_group)
        y = F.linear(x, self.weight, self.bias)
        return _GatherFromTP.apply(y, self.mp_group) if self.gather_output else y

    def extra_repr(self):
        return f"in={self.in_features}, out={self.out_features} (of {self.full_out_features}), tp={self.tp_world}"


class LinearAllreduce(nn.Module):
    """Row-parallel linear: weight ``[out, in/tp]``; partial results are all-reduced (or reduce-scattered)."""

    def __init__(self, module: nn.Linear, mp_group, mode="allreduce", grain=1, sub_params=None):
        super().__init__()
        self.mp_group, self.mode = mp_group, mode
        self.tp_world, self.tp_rank = _ws(mp_group), _rk(mp_group)
        w = shard_tensor(module.weight.data, 1, mp_group, sub_params, grain)
        self.weight = nn.Parameter(w, requires_grad=module.weight.requires_grad)
        self.weight.tensor_model_parallel = True
        self.weight.ds_tp_sharded = True
        self.weight.ds_tp_dim = 1
        self.bias = None
        if module.bias is not None:  # bias is replicated and added once after the reduction
            self.bias = nn.Parameter(module.bias.data.clone(), requires_grad=module.bias.requires_grad)
        self.in_features, self.out_features = self.weight.shape[1], self.weight.shape[0]

    def forward(self, x):
        y = F.linear(x, self.weight)
        if self.mode == "rs_ag":
            y = _ReduceScatterSeq.apply(y, self.mp_group)  # fused GEMM->RS kernel target
        else:
            y = _ReduceFromTP.apply(y, self.mp_group)
        return y + self.bias if self.bias is not None else y


class LmHeadLinearAllreduce(LinearAllreduce):
    """LM head sharded over the hidden (input) dimension; the input is sliced to the local shard."""

    def forward(self, x):
        sizes = _shard_sizes(x.shape[-1], self.tp_world)
        xs = x.split(sizes, dim=-1)[self.tp_rank] if x.shape[-1] != self.in_features else x
        return super().forward(xs)


class VocabParallelEmbedding(nn.Module):
    """Embedding sharded over the vocabulary; out-of-shard tokens contribute zeros, results are all-reduced."""

    def __init__(self, module: nn.Embedding, mp_group):
        super().__init__()
        self.mp_group = mp_group
        world, rank = _ws(mp_group), _rk(mp_group)
        sizes = _shard_sizes(module.weight.shape[0], world)
        self.start = sum(sizes[:rank])
        self.end = self.start + sizes[rank]
        self.weight = nn.Parameter(module.weight.data[self.start:self.end].clone())
        self.weight.tensor_model_parallel = True
        self.weight.ds_tp_sharded = True
        self.weight.ds_tp_dim = 0

    def forward(self, ids):
        mask = (ids < self.start) | (ids >= self.end)
        local = (ids - self.start).masked_fill(mask, 0)
        out = F.embedding(local, self.weight)
        out = out.masked_fill(mask.unsqueeze(-1), 0.0)
        return _ReduceFromTP.apply(out, self.mp_group)


# names of linears whose OUTPUT is reduced (row-parallel) in common architectures (reference auto_tp.py policy lists)
ROW_PARALLEL_PATTERNS = [r"\.o_proj$", r"\.out_proj$", r"\.down_proj$", r"\.dense_4h_to_h$", r"\.c_proj$", r"\.wo$",
                         r"\.w2$", r"attention\.dense$", r"\.fc2$", r"\.c_proj2$", r"self_attention\.dense$", r"\.fc_out$"]
COLUMN_PARALLEL_PATTERNS = [r"\.q_proj$", r"\.k_proj$", r"\.v_proj$", r"\.qkv_proj$", r"\.gate_proj$", r"\.up_proj$",
                            r"\.gate_up_proj$", r"\.query_key_value$", r"\.dense_h_to_4h$", r"\.c_attn$", r"\.c_fc$",
                            r"\.wq$", r"\.wk$", r"\.wv$", r"\.w1$", r"\.w3$", r"\.fc1$", r"\.Wqkv$", r"\.fc_in$"]


def _fused_sub_params(name, module, model):
    """Sub-matrix sizes of fused projections so q/k/v (and gate/up) are sharded consistently."""
    cfg = getattr(model, "cfg", None) or getattr(model, "config", None)
    out = module.weight.shape[0]
    if re.search(r"(qkv_proj|query_key_value|c_attn|Wqkv)$", name):
        if cfg is not None and hasattr(cfg, "num_attention_heads"):
            nh = cfg.num_attention_heads
            nkv = getattr(cfg, "num_key_value_heads", nh) or nh
            hd = getattr(cfg, "head_dim", None) or out // (nh + 2 * nkv)
            if (nh + 2 * nkv) * hd == out:
                return [nh * hd, nkv * hd, nkv * hd], hd
        return [out // 3] * 3, 1
    if re.search(r"gate_up_proj$", name):
        return [out // 2, out // 2], 1
    return None, 1


def _allreduce_grad(g, group):
    g = g.contiguous().clone()
    dist.all_reduce(g, group=group)
    return g


class AutoTP:
    """Walk a model and replace linears/embeddings with their tensor-parallel versions."""

    def __init__(self, model, mp_group, config=None, partition_config=None, mode="allreduce"):
        self.model, self.mp_group, self.config, self.mode = model, mp_group, config, mode
        self.partition_config = partition_config or {}
        # reference schema (module_inject/autotp_config.py): an AutoTPConfig, or a dict with layer_specs / preset
        self.specs = None
        from .autotp_config import AutoTPConfig, AutoTPPresets, merge_autotp_configs
        pc = self.partition_config
        if isinstance(pc, AutoTPConfig):
            self.specs, pc = pc, {}
        elif isinstance(pc, dict) and ("layer_specs" in pc or "use_default_specs" in pc):
            self.specs, pc = AutoTPConfig.from_dict(pc), {}
        preset = getattr(config, "preset_model", None)
        if preset:
            base = AutoTPPresets.get_preset(preset)
            if base is None:
                raise ValueError(f"unknown tensor_parallel.preset_model {preset!r}")
            self.specs = merge_autotp_configs(base, self.specs) if self.specs is not None else base
        self.partition_config = pc
        use_defaults = self.specs is None or self.specs.use_default_specs
        self.row = [re.compile(p) for p in (ROW_PARALLEL_PATTERNS if use_defaults else []) +
                    list(pc.get("row_parallel", []))]
        self.col = [re.compile(p) for p in (COLUMN_PARALLEL_PATTERNS if use_defaults else []) +
                    list(pc.get("column_parallel", []))]
        self.replaced = {}

    @staticmethod
    def from_hf_tp_plan(model):
        """Import a Hugging Face ``config.base_model_tp_plan`` (colwise / rowwise) as extra patterns."""
        plan = getattr(getattr(model, "config", None), "base_model_tp_plan", None) or {}
        cfg = {"row_parallel": [], "column_parallel": []}
        for pat, kind in plan.items():
            rx = pat.replace(".", r"\.").replace("*", r"\d+") + "$"
            (cfg["row_parallel"] if "row" in kind else cfg["column_parallel"]).append(rx)
        return cfg

    def _kind(self, name):
        if any(p.search(name) for p in self.row):
            return "row"
        if any(p.search(name) for p in self.col):
            return "col"
        return None

    def apply(self):
        world = _ws(self.mp_group)
        if world == 1:
            return self.model
        for name, module in list(self.model.named_modules()):
            if type(module).__name__ == "Conv1D" and getattr(module.weight, "dim", lambda: 0)() == 2:
                self._replace_conv1d(name, module)  # GPT-2 style ``x @ w`` projections (weight ``[in, out]``)
                continue
            if not isinstance(module, nn.Linear):
                continue
            spec = None
            if self.specs is not None:
                mtype = getattr(getattr(self.model, "config", None), "model_type", None)
                spec = self.specs.find_matching_spec(name + ".weight", mtype)
            if spec is not None:
                kind = spec.partition_type.value
                if kind == "skip":
                    if spec.grad_allreduce:  # replicated weight fed by rank-local activations: sum its gradient
                        group = self.mp_group
                        for p in module.parameters():
                            p.ds_tp_grad_allreduce = True
                            if p.requires_grad:
                                p.register_hook(lambda g, group=group: _allreduce_grad(g, group))
                    self.replaced[name] = "skip"
                    continue
            else:
                kind = self._kind("." + name)
            if kind is None:
                if self.specs is not None and self.specs.strict_mode:
                    raise ValueError(f"AutoTP strict_mode: no partition spec matches linear layer {name!r}")
                continue
            parent, attr = self._parent(name)
            if kind == "row":
                if spec is not None and getattr(spec, "split_input", False):
                    new = LmHeadLinearAllreduce(module, self.mp_group, mode=self.mode)
                else:
                    new = LinearAllreduce(module, self.mp_group, mode=self.mode)
            elif spec is not None:
                sub = spec.sub_param_sizes_for(module.weight.shape[0])
                grain = 1
                rows = None
                if spec.gather_output:
                    pass  # contiguous shards: the all-gather of the outputs restores the original column order
                elif sub is None:
                    rows = self._fused_qkv_rows(name, module, world)
                    if rows is None:
                        sub, grain = _fused_sub_params(name, module, self.model)
                new = LinearLayer(module, self.mp_group, sub_params=sub, gather_output=spec.gather_output,
                                  mode=self.mode, grain=grain, row_index=rows)
            else:
                rows = self._fused_qkv_rows(name, module, world)
                if rows is not None:
                    new = LinearLayer(module, self.mp_group, mode=self.mode, row_index=rows)
                else:
                    sub, grain = _fused_sub_params(name, module, self.model)
                    new = LinearLayer(module, self.mp_group, sub_params=sub, mode=self.mode, grain=grain)
            setattr(parent, attr, new)
            self.replaced[name] = kind
        self._fix_head_counts(world)
        cfg = getattr(self.model, "config", None)
        nh = getattr(cfg, "num_attention_heads", None) or getattr(cfg, "n_head", None) or getattr(cfg, "n_heads", None)
        if nh and self.replaced:
            from .auto_tp_model_utils import install_alibi_head_sharding
            install_alibi_head_sharding(self.model, self.mp_group, nh)
        log_dist(f"AutoTP: replaced {len(self.replaced)} linear layers (tp={world}, mode={self.mode})", [0])
        return self.model

    def _replace_conv1d(self, name, module):
        from .layers import Conv_LinearALlreduce, conv_LinearLayer
        kind = self._kind("." + name)
        if kind is None:
            return
        parent, attr = self._parent(name)
        if kind == "row":
            new = Conv_LinearALlreduce(module, self.mp_group, mode=self.mode)
        else:
            out = module.weight.shape[1]
            sub = [out // 3] * 3 if re.search(r"c_attn$", name) and out % 3 == 0 else None
            new = conv_LinearLayer(module, self.mp_group, sub_params=sub, mode=self.mode)
        setattr(parent, attr, new)
        self.replaced[name] = kind

    def _fused_qkv_rows(self, name, module, world):
        """Row index for a fused QKV projection whose enclosing block belongs to a known layout family (Bloom / Falcon /
        NeoX interleaved heads, CodeGen blocks, BigCode multi-query, Phi-3 / ChatGLM grouped KV); ``None`` for the
        plain ``[q; k; v]`` layout handled by sub-parameter sharding."""
        from . import fusedqkv_utils as Fq
        if not Fq.require_tp_fused_qkvw(name, world):
            return None
        parts, node, ftype = name.split("."), self.model, None
        for p in parts[:-1]:
            node = getattr(node, p)
            t = Fq.FUSED_QKV_TYPE_DICT.get(type(node).__name__)
            if t is None:
                hits = [k for k in Fq.FUSED_QKV_TYPE_DICT if k in type(node).__name__]
                t = Fq.FUSED_QKV_TYPE_DICT[max(hits, key=len)] if hits else None
            ftype = t or ftype
        if ftype in (None, "qwentype"):
            return None
        cfg = getattr(self.model, "config", None) or getattr(self.model, "cfg", None)
        nh = getattr(cfg, "num_attention_heads", None) or getattr(cfg, "n_head", None)
        nkv = getattr(cfg, "num_key_value_heads", None) or getattr(cfg, "multi_query_group_num", None) or \
            getattr(cfg, "num_kv_heads", None) or (1 if getattr(cfg, "multi_query", False) and
                                                   ftype == "bigcodetype" else nh)
        if ftype == "bloomtype" and getattr(cfg, "new_decoder_architecture", None) is False and \
                getattr(cfg, "multi_query", False):
            ftype, nkv = "bigcodetype", 1  # Falcon-7B multi-query layout [q heads; k; v]: shard q, replicate k / v
        n_embd = getattr(cfg, "hidden_size", None) or getattr(cfg, "n_embd", None)
        if ftype == "glmtype" and nkv != 2:
            return None  # plain [q; k; v]: the sub-parameter path shards it (also with grouped KV heads)
        Fq.set_num_attention_heads(nh), Fq.set_num_kv_heads(nkv), Fq.set_n_embd(n_embd)
        return Fq.fused_qkv_row_index(ftype, module.weight.shape[0], world, _rk(self.mp_group),
                                      hidden_cols=module.weight.shape[1])

    def _parent(self, name):
        parts = name.split(".")
        parent = self.model
        for p in parts[:-1]:
            parent = getattr(parent, p)
        return parent, parts[-1]

    def _fix_head_counts(self, world):
        """Attention modules keep per-rank head counts in attributes; divide them."""
        heads = ("nh", "nkv", "num_heads", "num_key_value_heads", "num_attention_heads", "n_head", "n_heads",
                 "num_kv_heads", "kv_n_heads", "num_attn_heads", "num_kv")
        # widths an attention module derives its reshapes / splits from (reference auto_tp.py ``update_mp_params``)
        widths = ("embed_dim", "all_head_size", "inner_dim", "split_size", "hidden_size", "d_model",
                  "hidden_size_per_partition")
        for m in self.model.modules():
            if isinstance(m, type(self.model)) or \
                    not any(isinstance(c, (LinearLayer, LinearAllreduce)) for c in m.children()):
                continue
            cols = [c for c in m.children() if isinstance(c, LinearLayer)]
            if cols and all(c.gather_output for c in cols):
                continue  # the projections' outputs are gathered: the module works on all heads
            is_attention = False
            for attr in heads:
                v = getattr(m, attr, None)
                if isinstance(v, int) and not isinstance(v, bool):
                    is_attention = True
                    if v % world == 0:
                        setattr(m, attr, v // world)
            if is_attention:
                for attr in widths:
                    v = getattr(m, attr, None)
                    if isinstance(v, int) and not isinstance(v, bool) and v % world == 0 and attr in vars(m):
                        setattr(m, attr, v // world)


def apply_autotp(model, tp_config, mp_group):
    """Entry point used by the engine (``tensor_parallel.autotp_size > 1``) and ``deepspeed.tp_model_init``."""
    pc = getattr(tp_config, "partition_config", None) or {}
    if not pc and not getattr(tp_config, "preset_model", None) and hasattr(model, "config"):
        # priority: explicit partition_config / preset > the model's Hugging Face tp_plan > name heuristics
        from .autotp_config import AutoTPConfig
        from .tp_plan_converter import TPPlanConverter, get_hf_tp_plan
        plan = get_hf_tp_plan(model)
        specs = TPPlanConverter.convert(plan) if plan else None
        if specs:
            pc = AutoTPConfig(tp_size=_ws(mp_group), layer_specs=specs, use_default_specs=False)
    mode = "allreduce"
    if getattr(tp_config, "fused_collectives", False) and isinstance(pc, dict) and \
            pc.get("sequence_parallel_activations", False):
        mode = "rs_ag"
    return AutoTP(model, mp_group, tp_config, pc, mode=mode).apply()


def universal_checkpoint_info(model):
    """UCP metadata for TP-sharded models (which params are row-parallel / replicated)."""
    row, rep, sub = [], [], []
    for name, p in model.named_parameters():
        if getattr(p, "ds_tp_sharded", False):
            if getattr(p, "ds_tp_dim", 0) == 1:
                row.append(re.escape(name) + "$")
        else:
            rep.append(re.escape(name) + "$")
    return {"universal_checkpoint_version": 0.2, "tp_replicated_parameter_patterns": rep,
            "parameter_with_row_parallelism_patterns": row}
