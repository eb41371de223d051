"""AutoEP: dropless token-choice expert parallelism (Mixtral / Qwen-MoE / DeepSeek style blocks).

Reference: ``module_inject/auto_ep_layer.py`` (``AutoEPMoELayer.forward`` :575-730: router -> stable argsort by
expert -> split plan via count all-to-all + host sync :95-133 -> ``_AllToAllV`` dispatch :681 -> permute to
expert-contiguous -> grouped GEMM SwiGLU (``moe/ep_experts.py:92-135`` ``torch._grouped_mm``) -> unpermute ->
``_AllToAllV`` combine :688 -> weighted sum :310), ``moe/ep_router.py:27``, ``auto_ep.py`` (HF block replacement),
``auto_ep_presets/*``.

Design: routing/sort/gather are index ops on the device; experts are stored stacked (``w13 [E_local, 2I, H]``,
``w2 [E_local, H, I]``) so the expert FFN is a grouped GEMM over expert-contiguous rows; expert parameters carry
``allreduce=False`` + ``group_name`` so ZeRO shards/reduces them over the expert-data-parallel group.
"""
import math
import re
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import comm as dist
from ..ops import fused
from ..ops.grouped_gemm import grouped_mm
from ..utils import groups
from ..utils.logging import log_dist


class _AllToAllV(torch.autograd.Function):
    """Variable-split all-to-all over rows; backward swaps the split lists."""

    @staticmethod
    def forward(ctx, group, x, in_splits, out_splits):
        ctx.group, ctx.in_splits, ctx.out_splits = group, in_splits, out_splits
        out = x.new_empty((sum(out_splits), ) + tuple(x.shape[1:]))
        dist.all_to_all_single(out, x.contiguous(), output_split_sizes=out_splits, input_split_sizes=in_splits,
                               group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        out = g.new_empty((sum(ctx.in_splits), ) + tuple(g.shape[1:]))
        dist.all_to_all_single(out, g.contiguous(), output_split_sizes=ctx.in_splits, input_split_sizes=ctx.out_splits,
                               group=ctx.group)
        return None, out, None, None


class TokenChoiceRouter(nn.Module):
    """Linear gate + softmax/sigmoid scoring + top-k (optionally renormalised), aux load-balancing loss."""

    def __init__(self, hidden, num_experts, top_k, score_func="softmax", norm_topk_prob=True, dtype=None, device=None,
                 aux_loss_coef=0.0):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(num_experts, hidden, dtype=dtype, device=device))
        nn.init.normal_(self.weight, std=0.02)
        self.num_experts, self.top_k, self.score_func, self.norm_topk_prob = num_experts, top_k, score_func, norm_topk_prob
        self.aux_loss_coef = aux_loss_coef
        self.l_aux = None

    def forward(self, x):
        logits = F.linear(x.float(), self.weight.float())
        scores = torch.softmax(logits, dim=-1) if self.score_func == "softmax" else torch.sigmoid(logits)
        top_w, top_i = torch.topk(scores, self.top_k, dim=-1)
        if self.norm_topk_prob:
            top_w = top_w / top_w.sum(dim=-1, keepdim=True).clamp(min=1e-20)
        if self.aux_loss_coef > 0 and self.training:
            me = scores.mean(dim=0)
            ce = torch.zeros_like(me).scatter_add_(0, top_i.reshape(-1), torch.ones_like(top_w.reshape(-1))) / top_i.numel()
            self.l_aux = self.aux_loss_coef * self.num_experts * torch.sum(me * ce)
        return top_w, top_i


class GroupedExperts(nn.Module):
    """``E_local`` SwiGLU experts with stacked weights, evaluated as grouped GEMMs over expert-sorted rows."""

    def __init__(self, num_local_experts, hidden, inter, dtype=None, device=None, group_name=None):
        super().__init__()
        self.w13 = nn.Parameter(torch.empty(num_local_experts, 2 * inter, hidden, dtype=dtype, device=device))
        self.w2 = nn.Parameter(torch.empty(num_local_experts, hidden, inter, dtype=dtype, device=device))
        nn.init.normal_(self.w13, std=0.02)
        nn.init.normal_(self.w2, std=0.02)
        for p in (self.w13, self.w2):
            p.allreduce = False
            p.group_name = group_name
        self.num_local_experts = num_local_experts

    def forward(self, x_sorted, counts):
        """x_sorted: [N, H] rows grouped by local expert (counts: python list of rows per expert)."""
        h = grouped_mm(x_sorted, self.w13, counts)
        h = fused.swiglu(h)
        return grouped_mm(h, self.w2, counts)


class AutoEPMoELayer(nn.Module):
    """Drop-in MoE block: ``forward(hidden [B,S,H]) -> [B,S,H]`` (HF blocks return ``(out, router_logits)``)."""

    def __init__(self, hidden, inter, num_experts, top_k, ep_size=1, dtype=None, device=None, score_func="softmax",
                 norm_topk_prob=True, return_router_logits=False, aux_loss_coef=0.0):
        super().__init__()
        assert num_experts % ep_size == 0
        self.hidden, self.num_experts, self.top_k, self.ep_size = hidden, num_experts, top_k, ep_size
        self.num_local_experts = num_experts // ep_size
        self.group_name = f"ep_size_{ep_size}"
        self.router = TokenChoiceRouter(hidden, num_experts, top_k, score_func, norm_topk_prob, torch.float32, device,
                                        aux_loss_coef)
        self.experts = GroupedExperts(self.num_local_experts, hidden, inter, dtype, device, self.group_name)
        self.return_router_logits = return_router_logits
        self.ep_group = None
        self.ep_rank = 0

    def _use_fused(self, x):
        import os
        # auto: fused NVLink dispatch/combine up to 4 expert-parallel ranks. Measured (profiles/moe_fused_dispatch_*.json):
        # 2.4 % faster than the NCCL all-to-all-v path at EP=2, 6 % slower at EP=8 (Mixtral 8x7B: 598 vs 563 ms/step; the
        # per-layer plan construction after the count sync and the 8-way READY/DONE handshakes cost more than they save).
        mode = os.environ.get("DSB200_FUSED_MOE", "auto")
        if mode == "0" or not x.is_cuda or (mode == "auto" and self.ep_size > 4):
            return False
        if not hasattr(self, "_fused_ok"):
            from ..moe import fused_dispatch as FD
            try:
                self._fused_ok = FD.supported(self.ep_group, self.hidden, self.top_k, x.dtype)
            except Exception:  # noqa: BLE001
                self._fused_ok = False
        return self._fused_ok

    def set_deepspeed_parallelism(self, use_data_before_expert_parallel_=False):
        if self.group_name not in groups._get_expert_parallel_group_dict():
            groups._create_expert_and_data_parallel(self.ep_size, use_data_before_expert_parallel_)
        self.ep_group = groups._get_expert_parallel_group(self.group_name)
        self.ep_rank = groups._get_expert_parallel_rank(self.group_name)

    def forward(self, hidden_states):
        shape = hidden_states.shape
        x = hidden_states.reshape(-1, shape[-1])
        T = x.shape[0]
        top_w, top_i = self.router(x)                           # [T, k]
        flat_e = top_i.reshape(-1)
        order = torch.argsort(flat_e, stable=True)              # token-slots sorted by expert
        tok = order // self.top_k
        routed = x.index_select(0, tok)                         # [T*k, H] expert-sorted
        counts = torch.bincount(flat_e, minlength=self.num_experts)
        if self.ep_size > 1 and self.ep_group is None:
            self.set_deepspeed_parallelism()
        if self.ep_size > 1 and self._use_fused(x):
            # ONE NVLink kernel per direction (csrc/moe_comm.cu): pull-dispatch into expert-contiguous rows, experts,
            # pull-combine with the router weights -- no index_select / all-to-all-v / permute / index_add passes
            from ..moe import fused_dispatch as FD
            comm = FD.FusedMoEComm.get(self.ep_group, self.hidden)
            plan = FD.build_plan(top_i, self.num_experts, self.ep_size, self.ep_rank, self.ep_group)
            expert_in = FD.fused_dispatch(x, comm, plan)
            expert_out = self.experts(expert_in, plan.local_counts)
            out = FD.fused_combine(expert_out, top_w, comm, plan).reshape(shape)
            if self.return_router_logits:
                return out, None
            return out
        if self.ep_size > 1:
            # split plan: counts [ep, E_local] exchanged so every rank knows what it receives (one host sync,
            # same as the reference; the fused NVLink dispatch removes it)
            recv_counts = torch.empty_like(counts)
            dist.all_to_all_single(recv_counts, counts, group=self.ep_group)
            c_send = counts.view(self.ep_size, self.num_local_experts)
            c_recv = recv_counts.view(self.ep_size, self.num_local_experts)
            in_splits = c_send.sum(dim=1).tolist()
            out_splits = c_recv.sum(dim=1).tolist()
            recv = _AllToAllV.apply(self.ep_group, routed, in_splits, out_splits)
            # received rows are grouped by source rank then by local expert -> regroup by local expert
            c_recv_l = c_recv.tolist()
            src_idx, exp_of_row = [], []
            offs = 0
            seg = []
            for r in range(self.ep_size):
                for e in range(self.num_local_experts):
                    seg.append((e, r, offs, c_recv_l[r][e]))
                    offs += c_recv_l[r][e]
            perm = torch.cat([torch.arange(o, o + n, device=x.device) for (e, r, o, n) in sorted(seg) if n > 0]) \
                if offs > 0 else torch.empty(0, dtype=torch.long, device=x.device)
            local_counts = [sum(c_recv_l[r][e] for r in range(self.ep_size)) for e in range(self.num_local_experts)]
            expert_in = recv.index_select(0, perm)
            expert_out = self.experts(expert_in, local_counts)
            back = torch.empty_like(expert_out)
            back.index_copy_(0, perm, expert_out)
            combined = _AllToAllV.apply(self.ep_group, back, out_splits, in_splits)
        else:
            combined = self.experts(routed, counts.tolist())
        w = top_w.reshape(-1).index_select(0, order).to(combined.dtype)
        out = torch.zeros_like(x)
        out.index_add_(0, tok, combined * w.unsqueeze(1))
        out = out.reshape(shape)
        if self.return_router_logits:
            return out, None
        return out

    @property
    def l_aux(self):
        return self.router.l_aux


# ---- replacing Hugging Face MoE blocks ---------------------------------------------------------------------
PRESETS = {
    "mixtral": dict(block=r"MixtralSparseMoeBlock", gate="gate", experts="experts", w1="w1", w2="w2", w3="w3"),
    "qwen3_moe": dict(block=r"Qwen3MoeSparseMoeBlock", gate="gate", experts="experts", w1="gate_proj", w2="down_proj",
                      w3="up_proj"),
    "qwen2_moe": dict(block=r"Qwen2MoeSparseMoeBlock", gate="gate", experts="experts", w1="gate_proj", w2="down_proj",
                      w3="up_proj"),
    "deepseek_v2": dict(block=r"DeepseekV2MoE", gate="gate", experts="experts", w1="gate_proj", w2="down_proj",
                        w3="up_proj"),
    "deepseek_v3": dict(block=r"DeepseekV3MoE", gate="gate", experts="experts", w1="gate_proj", w2="down_proj",
                        w3="up_proj"),
}


def _convert_block(block, preset, ep_size, top_k, ep_rank):
    gate = getattr(block, preset["gate"])
    experts = getattr(block, preset["experts"])
    E = len(experts)
    w1 = getattr(experts[0], preset["w1"]).weight
    inter, hidden = w1.shape
    layer = AutoEPMoELayer(hidden, inter, E, top_k or getattr(block, "top_k", 2), ep_size, dtype=w1.dtype,
                           device=w1.device, return_router_logits=True)
    with torch.no_grad():
        layer.router.weight.copy_(gate.weight.float())
        nl = E // ep_size
        for le in range(nl):
            ge = ep_rank * nl + le
            layer.experts.w13[le, :inter].copy_(getattr(experts[ge], preset["w1"]).weight)
            layer.experts.w13[le, inter:].copy_(getattr(experts[ge], preset["w3"]).weight)
            layer.experts.w2[le].copy_(getattr(experts[ge], preset["w2"]).weight)
    return layer


def apply_autoep(model, ep_config):
    """Replace every matching HF MoE block with :class:`AutoEPMoELayer` (``expert_parallel`` ds_config section)."""
    ep_size = int(getattr(ep_config, "autoep_size", 1) or 1)
    name = getattr(ep_config, "preset_model", None)
    presets = [PRESETS[name]] if name in PRESETS else list(PRESETS.values())
    if getattr(ep_config, "moe_layer_pattern", None):
        presets = [dict(block=ep_config.moe_layer_pattern, gate=ep_config.router_pattern or "gate",
                        experts=ep_config.expert_pattern or "experts", w1="w1", w2="w2", w3="w3")]
    if ep_size > 1:
        groups._create_expert_and_data_parallel(ep_size)
    ep_rank = groups._get_expert_parallel_rank(f"ep_size_{ep_size}") if ep_size > 1 else 0
    n = 0
    for mod_name, mod in list(model.named_modules()):
        for ps in presets:
            if re.search(ps["block"], type(mod).__name__) and hasattr(mod, ps["experts"]):
                parent = model
                parts = mod_name.split(".")
                for p in parts[:-1]:
                    parent = getattr(parent, p)
                new = _convert_block(mod, ps, ep_size, getattr(ep_config, "top_k", None), ep_rank)
                if ep_size > 1:
                    new.set_deepspeed_parallelism()
                setattr(parent, parts[-1], new)
                n += 1
                break
    log_dist(f"AutoEP: replaced {n} MoE blocks (ep_size={ep_size})", [0])
    return model
