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
from typing import NamedTuple, Optional

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


class RouterOutput(NamedTuple):
    top_scores: torch.Tensor  # [T, k] weights of the chosen experts (normalised / scaled)
    selected_experts: torch.Tensor  # [T, k] expert ids
    num_tokens_per_expert: torch.Tensor  # [E]
    logits: torch.Tensor  # [T, E] raw gate output


class _Gate(nn.Module):
    """Plain linear gate owned by the router when no model-provided gate module is adopted."""

    def __init__(self, dim, num_experts, bias, dtype, device):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(num_experts, dim, dtype=dtype, device=device))
        nn.init.normal_(self.weight, std=0.02)
        self.bias = nn.Parameter(torch.zeros(num_experts, dtype=dtype, device=device)) if bias else None


def _fire_forward_hooks(module, args, output):
    """Run the forward hooks registered on ``module`` (Transformers' router-logit recorders) for an output computed
    outside ``module.__call__``."""
    with_kwargs = getattr(module, "_forward_hooks_with_kwargs", {})
    for hid, hook in list(module._forward_hooks.items()):
        res = hook(module, args, {}, output) if with_kwargs.get(hid, False) else hook(module, args, output)
        if res is not None:
            output = res
    return output


class TokenChoiceTopKRouter(nn.Module):
    """Token-choice top-k routing: linear gate -> softmax / sigmoid scores -> optional group-limited (node-limited)
    masking -> top-k -> optional renormalisation and scaling; auxiliary load-balancing loss.

    Reference: ``moe/ep_router.py:27`` (``TokenChoiceTopKRouter``). ``gate`` may be the model's own router module
    (anything with a ``weight [E, H]``, optionally ``bias`` and the DeepSeek-V3 ``e_score_correction_bias``): it stays a
    submodule so checkpoints keep their names and hooks registered on it (Transformers records router logits that
    way) keep firing - ``hook_output`` says what they expect: ``"logits"`` or ``"topk_tuple"``
    (``(probabilities, top_scores, top_indices)``)."""

    def __init__(self, dim, num_experts, num_expert_groups=None, num_limited_groups=None, top_k=2,
                 score_func="softmax", route_norm=True, route_scale=1.0, gate_bias=False, group_score_func="top2_sum",
                 gate=None, dtype=None, device=None, aux_loss_coef=0.0, fp32_logits=None, hook_output="logits"):
        super().__init__()
        if score_func not in ("softmax", "sigmoid"):
            raise NotImplementedError(f"Unknown score function: {score_func}")
        if group_score_func not in ("max", "top2_sum"):
            raise NotImplementedError(f"Unknown group score function: {group_score_func}")
        if num_expert_groups is not None:
            if num_limited_groups is None:
                raise ValueError("num_limited_groups must be set when num_expert_groups is set")
            if num_limited_groups < 1 or num_limited_groups > num_expert_groups:
                raise ValueError(f"num_limited_groups must be in [1, {num_expert_groups}], got {num_limited_groups}")
            if num_experts % num_expert_groups != 0:
                raise ValueError(f"num_experts ({num_experts}) must be divisible by num_expert_groups "
                                 f"({num_expert_groups})")
            if group_score_func == "top2_sum" and num_experts // num_expert_groups < 2:
                raise ValueError("top2_sum group scoring needs at least 2 experts per group")
        self.gate = gate if gate is not None else _Gate(dim, num_experts, gate_bias, dtype, device)
        self.num_experts, self.top_k, self.score_func = num_experts, top_k, score_func
        self.num_expert_groups, self.num_limited_groups = num_expert_groups, num_limited_groups
        self.group_score_func = group_score_func
        self.route_norm, self.route_scale = route_norm, float(route_scale)
        self.aux_loss_coef = aux_loss_coef
        self.fp32_logits = fp32_logits
        self.hook_output = hook_output
        self.l_aux = None
        self.tp_group = None  # set by the folded (AutoEP over AutoTP) layer: the weight saw a token slice only

    # the gate weight under its historical name (round-1 checkpoints / tests use ``router.weight``)
    @property
    def weight(self):
        return self.gate.weight

    @property
    def norm_topk_prob(self):
        return self.route_norm

    @property
    def e_score_correction_bias(self):
        return getattr(self.gate, "e_score_correction_bias", None)

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        old = prefix + "weight"
        if old in state_dict and prefix + "gate.weight" not in state_dict:
            state_dict[prefix + "gate.weight"] = state_dict.pop(old)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)

    def _group_limited(self, scores):
        """Keep the experts of the ``num_limited_groups`` best groups, ``-inf`` elsewhere."""
        T = scores.shape[0]
        per = self.num_experts // self.num_expert_groups
        grouped = scores.view(T, self.num_expert_groups, per)
        if self.group_score_func == "max":
            group_scores = grouped.max(dim=-1).values
        else:
            group_scores = grouped.topk(2, dim=-1).values.sum(dim=-1)
        keep = torch.topk(group_scores, self.num_limited_groups, dim=-1, sorted=False).indices
        mask = torch.zeros_like(group_scores, dtype=torch.bool).scatter_(1, keep, True)
        mask = mask.unsqueeze(-1).expand(T, self.num_expert_groups, per).reshape(T, self.num_experts)
        return scores.masked_fill(~mask, float("-inf"))

    def logits(self, x):
        w, b = self.gate.weight, getattr(self.gate, "bias", None)
        if self.tp_group is not None:
            from .auto_ep_folding import sum_grad_over_lanes
            w = sum_grad_over_lanes(w, self.tp_group)
            b = sum_grad_over_lanes(b, self.tp_group) if b is not None else None
        fp32 = self.fp32_logits if self.fp32_logits is not None else w.dtype == torch.float32
        if fp32:
            return F.linear(x.float(), w.float(), b.float() if b is not None else None)
        return F.linear(x.to(w.dtype), w, b).float()

    def forward(self, x, expert_bias=None, valid_rows=None):
        """x: ``[T, H]``. ``expert_bias [E]``: added to the scores used for *choosing* experts only (the weights come
        from the unbiased scores), on top of a trained ``e_score_correction_bias`` of the gate."""
        logits = self.logits(x)
        scores = torch.softmax(logits, dim=-1) if self.score_func == "softmax" else torch.sigmoid(logits)
        choice = scores
        corr = self.e_score_correction_bias
        if corr is not None:
            choice = choice + corr.to(choice.dtype)
        if expert_bias is not None:
            choice = choice + expert_bias.to(choice.dtype)
        if self.num_expert_groups is not None:
            choice = self._group_limited(choice)
        top_i = torch.topk(choice, self.top_k, dim=-1).indices
        top_w = scores.gather(1, top_i)
        if self.route_norm:
            top_w = top_w / top_w.sum(dim=-1, keepdim=True).clamp(min=1e-20)
        if self.route_scale != 1.0:
            top_w = top_w * self.route_scale
        counts = torch.bincount(top_i.reshape(-1) if valid_rows is None else top_i[:valid_rows].reshape(-1),
                                minlength=self.num_experts)
        if self.aux_loss_coef > 0 and self.training:
            sc = scores if valid_rows is None else scores[:valid_rows]
            me = sc.mean(dim=0)
            ce = counts.to(me.dtype) / max(1, int(sc.shape[0]) * self.top_k)
            self.l_aux = self.aux_loss_coef * self.num_experts * torch.sum(me * ce)
        if getattr(self.gate, "_forward_hooks", None):
            out = logits if self.hook_output == "logits" else (scores, top_w, top_i)
            _fire_forward_hooks(self.gate, (x, ), out)
        return RouterOutput(top_w, top_i, counts, logits)


class TokenChoiceRouter(TokenChoiceTopKRouter):
    """Round-1 constructor signature."""

    def __init__(self, hidden, num_experts, top_k, score_func="softmax", norm_topk_prob=True, dtype=None, device=None,
                 aux_loss_coef=0.0, **kw):
        super().__init__(hidden, num_experts, top_k=top_k, score_func=score_func, route_norm=norm_topk_prob,
                         dtype=dtype, device=device, aux_loss_coef=aux_loss_coef, **kw)


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


class SplitPlan(NamedTuple):
    """What one rank sends / receives in the expert dispatch (reference ``auto_ep_layer.py:41``)."""
    input_splits: list  # rows sent to every expert-parallel rank
    output_splits: list  # rows received from every rank
    local_counts: list  # rows per local expert after regrouping
    perm: torch.Tensor  # received row order -> expert-contiguous order


def count_tokens_per_expert(selected_experts, num_experts):
    """Assignments per expert (reference ``moe/ep_count.py``)."""
    return torch.bincount(selected_experts.reshape(-1), minlength=num_experts)


def compute_split_plan(counts, ep_group, ep_size, num_local_experts):
    """Exchange the per-expert counts so that every rank knows what it receives (one host read, as in the
    reference; the fused NVLink dispatch of ``moe/fused_dispatch.py`` avoids it)."""
    recv_counts = torch.empty_like(counts)
    dist.all_to_all_single(recv_counts, counts, group=ep_group)
    c_send = counts.view(ep_size, num_local_experts)
    c_recv = recv_counts.view(ep_size, num_local_experts)
    out_splits = c_recv.sum(dim=1).tolist()  # the one host read: all-to-all-v needs the sizes on the host
    local_counts = c_recv.sum(dim=0).tolist()
    # received rows are grouped by source rank then by local expert -> regroup by local expert (index math on the device)
    perm = regroup_indices(c_recv, sum(out_splits))
    return SplitPlan(c_send.sum(dim=1).tolist(), out_splits, local_counts, perm)


def regroup_indices(counts_rank_expert, total=None):
    """For rows laid out ``[rank][local expert]`` with the given counts ``[ranks, experts]``: the gather index that
    lists them ``[local expert][rank]`` (expert-contiguous). Vectorised - no loop over the segments."""
    R, E = counts_rank_expert.shape
    flat = counts_rank_expert.reshape(-1).long()
    src_start = (torch.cumsum(flat, 0) - flat).view(R, E).t().reshape(-1)  # start of segment (r, e), listed (e, r)
    by_expert = counts_rank_expert.t().reshape(-1).long()
    dst_start = torch.cumsum(by_expert, 0) - by_expert
    total = int(flat.sum()) if total is None else int(total)
    if total == 0:
        return torch.empty(0, dtype=torch.long, device=counts_rank_expert.device)
    seg = torch.repeat_interleave(torch.arange(R * E, device=flat.device), by_expert, output_size=total)
    return src_start[seg] + (torch.arange(total, device=flat.device) - dst_start[seg])


def generate_permute_indices(tokens_per_expert_group, experts_per_rank, num_ranks, max_len, alignment, use_cpu=False):
    """Aligned variant (reference ``moe/ep_kernels.py:190``, a Triton kernel there): every local expert's block is padded
    to a multiple of ``alignment`` (and at least ``alignment``) so that grouped GEMM tiles never straddle experts.
    Returns ``(permuted_indices [max_len] with -1 in the padding, m_sizes int32 [experts], m_offsets)``; everything
    stays on the device."""
    counts = tokens_per_expert_group.view(num_ranks, experts_per_rank).long()
    dev = counts.device
    total_per_expert = counts.sum(0).clamp_min(alignment)
    m_sizes = ((total_per_expert + alignment - 1) // alignment * alignment).to(torch.int32)
    m_offsets = torch.cumsum(m_sizes, 0)
    write_start = (m_offsets - m_sizes).long()  # first row of every expert's padded block
    flat = counts.reshape(-1)
    src_start = (torch.cumsum(flat, 0) - flat).view(num_ranks, experts_per_rank)
    by_expert = counts.t()  # [experts, ranks]
    within_expert = torch.cumsum(by_expert, 1) - by_expert  # offset of rank r's rows inside expert e's block
    out = torch.full((max_len, ), -1, dtype=torch.int32, device=dev)
    n = int(flat.sum())
    if n:
        seg = torch.repeat_interleave(torch.arange(experts_per_rank * num_ranks, device=dev), by_expert.reshape(-1),
                                      output_size=n)
        e, r = seg // num_ranks, seg % num_ranks
        seg_first = torch.cumsum(by_expert.reshape(-1), 0) - by_expert.reshape(-1)
        i = torch.arange(n, device=dev) - seg_first[seg]
        dst = write_start[e] + within_expert[e, r] + i
        keep = dst < max_len
        out[dst[keep]] = (src_start[r, e] + i)[keep].to(torch.int32)
    return out, m_sizes, m_offsets.to(torch.int32)




def compute_split_plan_from_expert_indices(selected_experts, num_experts, ep_group, ep_size):
    return compute_split_plan(count_tokens_per_expert(selected_experts, num_experts), ep_group, ep_size,
                              num_experts // ep_size)


def permute_by_local_expert(received, plan):
    """Rows as they arrived (by source rank, then expert) -> grouped by local expert."""
    return received.index_select(0, plan.perm)


def unpermute_by_local_expert(expert_out, plan):
    """Inverse of :func:`permute_by_local_expert`."""
    back = torch.empty_like(expert_out)
    back.index_copy_(0, plan.perm, expert_out)
    return back


def resolve_score_apply_mode(config_mode, preset_mode):
    return preset_mode if config_mode in (None, "auto") else config_mode


def resolve_combine_impl(combine_impl):
    """``auto`` and ``weighted_sum`` are the index-add combine of this layer; ``legacy_bmm`` (a dense one-hot batched
    matmul in the reference) is accepted and served by the same code."""
    if combine_impl not in ("auto", "weighted_sum", "legacy_bmm"):
        raise ValueError(f"combine_impl must be one of ('auto', 'weighted_sum', 'legacy_bmm'), got '{combine_impl}'")
    return "weighted_sum"


def apply_scores_before_experts_if_enabled(routed, scores, score_apply):
    return routed * scores.unsqueeze(1).to(routed.dtype) if score_apply == "pre" else routed


def combine_from_routed(routed_out, token_of_row, scores, num_tokens, score_apply="post"):
    """``out[t] = sum over the rows of token t`` (weighted by the router score unless it was applied before)."""
    if score_apply == "post":
        routed_out = routed_out * scores.unsqueeze(1).to(routed_out.dtype)
    out = routed_out.new_zeros(num_tokens, routed_out.shape[1])
    out.index_add_(0, token_of_row, routed_out)
    return out


class AutoEPMoELayer(nn.Module):
    """Drop-in MoE block: ``forward(hidden [B,S,H]) -> [B,S,H]`` (or ``(out, router_logits)`` when the block it
    replaces returned that pair).

    ``score_apply``: ``"post"`` weights the expert outputs in the combine, ``"pre"`` scales the rows going into the
    experts. ``shared_experts`` (with an optional sigmoid ``shared_experts_gate``) run on every token next to the
    routed experts. ``tp_group``: AutoEP folded over AutoTP - this lane routes only its slice of the replicated
    tokens (``parallel/auto_ep_folding.py``)."""

    def __init__(self, hidden, inter, num_experts, top_k, ep_size=1, dtype=None, device=None, score_func="softmax",
                 norm_topk_prob=True, return_router_logits=False, aux_loss_coef=0.0, *, route_scale=1.0,
                 score_apply="post", num_expert_groups=None, num_limited_groups=None, group_score_func="top2_sum",
                 gate_bias=False, gate=None, shared_experts=None, shared_experts_gate=None, output_shape="batched",
                 fp32_logits=None, hook_output="logits", tp_group=None):
        super().__init__()
        assert num_experts % ep_size == 0, f"num_experts={num_experts} is not divisible by ep_size={ep_size}"
        assert score_apply in ("pre", "post")
        self.hidden, self.num_experts, self.top_k, self.ep_size = hidden, num_experts, top_k, ep_size
        self.num_local_experts = num_experts // ep_size
        self.group_name = f"ep_size_{ep_size}"
        self.router = TokenChoiceTopKRouter(hidden, num_experts, num_expert_groups, num_limited_groups, top_k,
                                            score_func, norm_topk_prob, route_scale, gate_bias, group_score_func,
                                            gate=gate, dtype=torch.float32, device=device,
                                            aux_loss_coef=aux_loss_coef, fp32_logits=fp32_logits,
                                            hook_output=hook_output)
        self.experts = GroupedExperts(self.num_local_experts, hidden, inter, dtype, device, self.group_name)
        self.shared_experts = shared_experts
        self.shared_experts_gate = shared_experts_gate
        self.score_apply = score_apply
        self.output_shape = output_shape
        self.return_router_logits = return_router_logits
        self.ep_group = None
        self.ep_rank = 0
        self.tp_group = None
        self.set_folding(tp_group)

    def set_folding(self, tp_group):
        """Route a 1/tp slice of the (TP-replicated) tokens on this lane; ``None`` turns folding off."""
        self.tp_group = tp_group if tp_group is not None and dist.get_world_size(group=tp_group) > 1 else None
        self.router.tp_group = self.tp_group

    def _use_fused(self, x):
        import os
        # auto: fused NVLink dispatch/combine up to 4 expert-parallel ranks. Measured (profiles/moe_fused_dispatch_*.json):
        # 2.4 % faster than the NCCL all-to-all-v path at EP=2, 6 % slower at EP=8 (Mixtral 8x7B: 598 vs 563 ms/step; the
        # per-layer plan construction after the count sync and the 8-way READY/DONE handshakes cost more than they save).
        mode = os.environ.get("DSB200_FUSED_MOE", "auto")
        if mode == "0" or not x.is_cuda or (mode == "auto" and self.ep_size > 4) or self.score_apply == "pre":
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

    def _routed(self, x, valid_rows=None):
        """Router + dispatch + experts + combine on ``x [T, H]``; returns (``[T, H]``, raw router logits)."""
        r = self.router(x, valid_rows=valid_rows)
        top_w, top_i = r.top_scores, r.selected_experts
        flat_e = top_i.reshape(-1)
        order = torch.argsort(flat_e, stable=True)              # token-slots sorted by expert
        tok = order // self.top_k
        counts = count_tokens_per_expert(flat_e, self.num_experts)
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
            return FD.fused_combine(expert_out, top_w, comm, plan), r.logits
        routed = x.index_select(0, tok)                         # [T*k, H] expert-sorted
        w = top_w.reshape(-1).index_select(0, order).to(x.dtype)
        routed = apply_scores_before_experts_if_enabled(routed, w, self.score_apply)
        if self.ep_size > 1:
            plan = compute_split_plan(counts, self.ep_group, self.ep_size, self.num_local_experts)
            recv = _AllToAllV.apply(self.ep_group, routed, plan.input_splits, plan.output_splits)
            expert_out = self.experts(permute_by_local_expert(recv, plan), plan.local_counts)
            combined = _AllToAllV.apply(self.ep_group, unpermute_by_local_expert(expert_out, plan),
                                        plan.output_splits, plan.input_splits)
        else:
            combined = self.experts(routed, counts.tolist())
        return combine_from_routed(combined, tok, w, x.shape[0], self.score_apply), r.logits

    def forward(self, hidden_states):
        shape = hidden_states.shape
        x = hidden_states.reshape(-1, shape[-1])
        if self.tp_group is not None:
            from .auto_ep_folding import partition_tokens, restore_tokens
            part, real = partition_tokens(x, self.tp_group)
            out, logits = self._routed(part, valid_rows=real)
            out = restore_tokens(out, x.shape[0], self.tp_group)
            if self.return_router_logits:
                logits = restore_tokens(logits, x.shape[0], self.tp_group)
        else:
            out, logits = self._routed(x)
        if self.shared_experts is not None:
            sh = self.shared_experts(x)
            if self.shared_experts_gate is not None:
                sh = torch.sigmoid(self.shared_experts_gate(x)) * sh
            out = out + sh
        if self.output_shape == "batched":
            out = out.reshape(shape)
        if self.return_router_logits:
            return out, logits
        return out

    @property
    def l_aux(self):
        return self.router.l_aux


# ---- replacing a model's MoE blocks -----------------------------------------------------------------------
def _param_of(obj, name):
    p = getattr(obj, name, None)
    if p is None:
        return None
    return p.weight if isinstance(p, nn.Module) and hasattr(p, "weight") else p


def _route_scale(config, model_config):
    v = config.routed_scaling_factor
    if v == "auto":
        hf = getattr(model_config, "routed_scaling_factor", None) if model_config is not None else None
        v = hf if hf is not None and config.route_scale == 1.0 else config.route_scale
    if isinstance(v, bool) or not isinstance(v, (int, float)) or not math.isfinite(float(v)):
        raise ValueError(f"routed_scaling_factor / route_scale must be a finite number, got {v!r}")
    return float(v)


def _name_matches(pattern, name, module):
    """``pattern`` against the module path (also without a leading ``model.`` - a bare backbone has none - and as a
    path suffix, for backbones nested in a wrapper) or against the class name."""
    for p in {pattern, pattern[len(r"model\."):] if pattern.startswith(r"model\.") else pattern}:
        if re.search(r"(?:^|\.)" + p + "$", name):
            return True
    return re.fullmatch(pattern, type(module).__name__) is not None


class AutoEP:
    """Finds the MoE blocks of a model (``ep_parser``) and swaps each for an :class:`AutoEPMoELayer` holding this
    rank's experts (``replace_moe_layer(s)``).

    Reference: ``module_inject/auto_ep.py`` (``AutoEP.ep_parser`` :283, ``replace_moe_layer`` :528). Expert weights are
    read from either layout - fused 3-D parameters (``gate_up_proj [E, 2I, H]`` / ``[E, H, 2I]`` classified by
    ``moe/fused_expert_layout.py``, or separate gate / up tensors) or a ``ModuleList`` of MLPs - and written into the
    stacked ``w13`` / ``w2`` of the grouped-GEMM experts; the model's router module and shared experts are adopted as
    they are."""

    def __init__(self, model, config):
        from . import auto_ep_presets as P
        self.model = model
        self.config = config if isinstance(config, P.AutoEPConfig) else P.parse_autoep_config(config)
        self.model_config = getattr(model, "config", None)
        self.specs = None

    # -- detection ------------------------------------------------------------------------------------------
    def _spec_for(self, name, module, preset_name, preset):
        from . import auto_ep_presets as P
        from ..moe.fused_expert_layout import classify_fused_gate_up_layout
        cfg, mc = self.config, self.model_config
        router = getattr(module, preset.router_pattern, None)
        experts = getattr(module, preset.experts_pattern, None)
        if router is None or experts is None or not hasattr(router, "weight"):
            return None
        adapter = P.get_preset_adapter(preset.preset_adapter)
        preset = adapter.resolve_expert_layout(experts, preset)
        storage = "module_list" if isinstance(experts, nn.ModuleList) else "fused_3d"
        if storage == "fused_3d" and preset.expert_storage == "module_list":
            return None  # (a ModuleList with separate w1/w2/w3 names is accepted whatever the preset's default says)
        if storage == "module_list":
            if len(experts) == 0 or preset.expert_w3 is None:
                return None
            w1 = _param_of(experts[0], preset.expert_w1)
            w2 = _param_of(experts[0], preset.expert_w2)
            if w1 is None or w2 is None or _param_of(experts[0], preset.expert_w3) is None:
                return None
            n_exp, (ffn, hidden) = len(experts), tuple(w1.shape)
        else:
            w1, w2 = _param_of(experts, preset.expert_w1), _param_of(experts, preset.expert_w2)
            if w1 is None or w2 is None or w1.dim() != 3 or w2.dim() != 3:
                return None
            n_exp = w1.shape[0]
            if preset.expert_w3 is None:
                lay = classify_fused_gate_up_layout(tuple(w1.shape), tuple(w2.shape))
                if lay is None:
                    return None
                hidden, ffn = lay.hidden_size, lay.ffn_hidden_size
            else:
                w3 = _param_of(experts, preset.expert_w3)
                if w3 is None or w3.shape != w1.shape:
                    return None
                if tuple(w2.shape[1:]) == (w1.shape[2], w1.shape[1]):
                    ffn, hidden = w1.shape[1], w1.shape[2]
                else:
                    hidden, ffn = w1.shape[1], w1.shape[2]
        cfg_n = getattr(mc, preset.num_experts_attr, None) if mc is not None else None
        if cfg_n is not None and int(cfg_n) != n_exp:
            raise ValueError(f"AutoEP: layer '{name}' holds {n_exp} experts but config.{preset.num_experts_attr}="
                             f"{cfg_n}")
        if isinstance(cfg.top_k, int) and not isinstance(cfg.top_k, bool):
            top_k = cfg.top_k
        else:
            top_k = getattr(mc, preset.top_k_attr, None) if mc is not None else None
            top_k = top_k if top_k is not None else getattr(module, "top_k", getattr(router, "top_k", None))
        if top_k is None:
            raise ValueError(f"AutoEP: cannot determine top_k for layer '{name}' (set expert_parallel.top_k)")
        gr = adapter.resolve_group_routing(cfg, mc)
        contract = adapter.adjust_forward_contract(P.ForwardContract(), module, storage)
        if cfg.return_router_logits is not None:
            contract = P.ForwardContract(**{**contract.__dict__, "return_router_logits": cfg.return_router_logits})
        has_shared = bool(preset.has_shared_experts and preset.shared_experts_pattern
                          and getattr(module, preset.shared_experts_pattern, None) is not None)
        gate_name = preset.shared_experts_gate_pattern if has_shared else ""
        if gate_name and getattr(module, gate_name, None) is None:
            gate_name = ""
        return P.MoELayerSpec(
            moe_module_name=name, model_family=preset_name, router_name=preset.router_pattern,
            experts_name=preset.experts_pattern, expert_storage=storage, expert_w1_name=preset.expert_w1,
            expert_w2_name=preset.expert_w2, expert_w3_name=preset.expert_w3, num_experts=n_exp, top_k=int(top_k),
            hidden_size=int(hidden), ffn_hidden_size=int(ffn),
            score_func=cfg.score_func if cfg.score_func != "auto" else preset.score_func,
            score_apply=cfg.score_apply if cfg.score_apply != "auto" else preset.score_apply,
            route_norm=adapter.resolve_route_norm(cfg, preset, mc),
            gate_bias=getattr(router, "bias", None) is not None,
            return_router_logits=contract.return_router_logits, router_logits_capture_target=contract.capture_target,
            router_logits_capture_index=contract.capture_index,
            router_logits_capture_layer_name=contract.capture_layer_name, has_shared_experts=has_shared,
            shared_experts_name=preset.shared_experts_pattern if has_shared else "", shared_experts_gate_name=gate_name,
            route_scale=_route_scale(cfg, mc), num_expert_groups=gr.num_expert_groups,
            num_limited_groups=gr.num_limited_groups, group_score_func=gr.group_score_func,
            supports_expert_bias=preset.supports_expert_bias,
            unsupported_router_bias_names=preset.unsupported_router_bias_names, preset_adapter=preset.preset_adapter,
            router_logits_capture_mode=contract.router_logits_capture_mode,
            moe_output_shape=contract.moe_output_shape)

    def ep_parser(self):
        """Detected :class:`MoELayerSpec` list (validated against ``autoep_size`` and the routing-group settings)."""
        from . import auto_ep_presets as P
        candidates = P.resolve_preset_candidates(self.config, self.model_config)
        if len(candidates) == 1:
            self.config = P.resolve_autoep_config_defaults(self.config, candidates[0][0])
        specs, seen = [], set()
        for name, module in self.model.named_modules():
            if isinstance(module, AutoEPMoELayer) or name in seen:
                continue
            for preset_name, preset in candidates:
                if not _name_matches(preset.moe_layer_pattern, name, module):
                    continue
                spec = self._spec_for(name, module, preset_name, preset)
                if spec is not None:
                    specs.append(spec)
                    seen.add(name)
                    break
        if not specs and (self.config.preset_model is not None or self.config.moe_layer_pattern):
            tried = ", ".join(f"{n} ({p.moe_layer_pattern})" for n, p in candidates)
            raise ValueError(f"AutoEP: no MoE layers detected with preset(s) {tried}. Check moe_layer_pattern / "
                             "router_pattern / expert_pattern against the model's module names.")
        P.validate_autoep_post_detection(self.config, specs)
        self.specs = specs
        return specs

    # -- replacement ------------------------------------------------------------------------------------------
    def _copy_experts(self, spec, experts, layer, ep_rank):
        from ..moe.fused_expert_layout import classify_fused_gate_up_layout
        nl, inter = layer.num_local_experts, spec.ffn_hidden_size
        lo = ep_rank * nl
        with torch.no_grad():
            if spec.expert_storage == "module_list":
                for le in range(nl):
                    e = experts[lo + le]
                    layer.experts.w13[le, :inter].copy_(_param_of(e, spec.expert_w1_name))
                    layer.experts.w13[le, inter:].copy_(_param_of(e, spec.expert_w3_name))
                    layer.experts.w2[le].copy_(_param_of(e, spec.expert_w2_name))
                return
            w1 = _param_of(experts, spec.expert_w1_name)[lo:lo + nl]
            w2 = _param_of(experts, spec.expert_w2_name)[lo:lo + nl]
            if spec.expert_w3_name is None:
                lay = classify_fused_gate_up_layout(tuple(_param_of(experts, spec.expert_w1_name).shape),
                                                    tuple(_param_of(experts, spec.expert_w2_name).shape))
                layer.experts.w13.copy_(w1.transpose(1, 2) if lay.needs_transpose else w1)
                layer.experts.w2.copy_(w2.transpose(1, 2) if lay.needs_transpose else w2)
            else:
                w3 = _param_of(experts, spec.expert_w3_name)[lo:lo + nl]
                t = tuple(w1.shape[1:]) != (inter, spec.hidden_size)
                layer.experts.w13[:, :inter].copy_(w1.transpose(1, 2) if t else w1)
                layer.experts.w13[:, inter:].copy_(w3.transpose(1, 2) if t else w3)
                layer.experts.w2.copy_(w2.transpose(1, 2) if t else w2)

    def replace_moe_layer(self, spec, ep_size=None, ep_rank=None, tp_group=None):
        ep_size = int(ep_size if ep_size is not None else self.config.autoep_size)
        if ep_rank is None:
            ep_rank = groups._get_expert_parallel_rank(f"ep_size_{ep_size}") if ep_size > 1 else 0
        parent, leaf = self.model, spec.moe_module_name
        if "." in leaf:
            parent = self.model.get_submodule(leaf.rsplit(".", 1)[0])
            leaf = leaf.rsplit(".", 1)[1]
        block = getattr(parent, leaf)
        router, experts = getattr(block, spec.router_name), getattr(block, spec.experts_name)
        ref = _param_of(experts[0], spec.expert_w1_name) if spec.expert_storage == "module_list" else \
            _param_of(experts, spec.expert_w1_name)
        unsupported = [n for n in spec.unsupported_router_bias_names if getattr(router, n, None) is not None]
        if unsupported:
            raise ValueError(f"AutoEP: router of '{spec.moe_module_name}' carries unsupported bias term(s) "
                             f"{unsupported}")
        layer = AutoEPMoELayer(
            spec.hidden_size, spec.ffn_hidden_size, spec.num_experts, spec.top_k, ep_size, dtype=ref.dtype,
            device=ref.device, score_func=spec.score_func, norm_topk_prob=spec.route_norm,
            return_router_logits=spec.return_router_logits, route_scale=spec.route_scale,
            score_apply=spec.score_apply, num_expert_groups=spec.num_expert_groups,
            num_limited_groups=spec.num_limited_groups, group_score_func=spec.group_score_func, gate=router,
            shared_experts=getattr(block, spec.shared_experts_name) if spec.has_shared_experts else None,
            shared_experts_gate=getattr(block, spec.shared_experts_gate_name) if spec.shared_experts_gate_name else None,
            output_shape=spec.moe_output_shape, fp32_logits=True if spec.model_family.startswith("deepseek") else None,
            hook_output="topk_tuple" if spec.preset_adapter in ("router_hook", "qwen3_5_moe") and
            spec.router_logits_capture_target == "router" else "logits", tp_group=tp_group)
        self._copy_experts(spec, experts, layer, ep_rank)
        g13, g2 = repack_expert_requires_grad_flags(experts, spec)
        layer.experts.w13.requires_grad_(g13)
        layer.experts.w2.requires_grad_(g2)
        layer.train(block.training)
        layer._autoep_spec = spec
        setattr(parent, leaf, layer)
        return layer

    def replace_moe_layers(self, specs=None, tp_group=None):
        specs = specs if specs is not None else (self.specs if self.specs is not None else self.ep_parser())
        ep_size = int(self.config.autoep_size)
        if ep_size > 1:
            groups._create_expert_and_data_parallel(ep_size)
        layers = []
        for spec in specs:
            layer = self.replace_moe_layer(spec, ep_size, tp_group=tp_group)
            if ep_size > 1:
                layer.set_deepspeed_parallelism()
            layers.append(layer)
        log_dist(f"AutoEP: replaced {len(layers)} MoE blocks (ep_size={ep_size}"
                 f"{', folded over TP' if tp_group is not None else ''})", [0])
        return layers


def _source_params(experts_source, spec, name):
    if spec.expert_storage == "module_list":
        return [_param_of(e, name) for e in experts_source]
    return [_param_of(experts_source, name)]


def repack_expert_requires_grad_flags(experts_source, spec):
    """(gate/up trainable, down trainable): the stacked tensors are trainable when any of their source weights is."""
    up = _source_params(experts_source, spec, spec.expert_w1_name)
    if spec.expert_w3_name is not None:
        up = up + _source_params(experts_source, spec, spec.expert_w3_name)
    down = _source_params(experts_source, spec, spec.expert_w2_name)
    return any(p.requires_grad for p in up), any(p.requires_grad for p in down)


def repack_expert_weights(experts_source, spec, ep_rank, ep_size):
    """This rank's experts in the grouped layout, from either storage of the model (reference ``moe/ep_repack.py:39``):
    ``(w1 gate [E_local, I, H], w2 down [E_local, H, I], w3 up [E_local, I, H])``."""
    from ..moe.fused_expert_layout import classify_fused_gate_up_layout
    nl, inter = spec.num_experts // ep_size, spec.ffn_hidden_size
    lo = ep_rank * nl
    if spec.expert_storage == "module_list":
        pick = lambda name: torch.stack([_param_of(experts_source[lo + i], name).detach() for i in range(nl)])  # noqa: E731
        return pick(spec.expert_w1_name), pick(spec.expert_w2_name), pick(spec.expert_w3_name)
    if spec.expert_storage != "fused_3d":
        raise ValueError(f"Unknown expert_storage type: {spec.expert_storage}")
    w1 = _param_of(experts_source, spec.expert_w1_name).detach()[lo:lo + nl]
    w2 = _param_of(experts_source, spec.expert_w2_name).detach()[lo:lo + nl]
    if spec.expert_w3_name is None:
        lay = classify_fused_gate_up_layout(tuple(_param_of(experts_source, spec.expert_w1_name).shape),
                                            tuple(_param_of(experts_source, spec.expert_w2_name).shape))
        if lay is None:
            raise ValueError("fused gate/up expert tensors have an unrecognised layout")
        if lay.needs_transpose:
            w1, w2 = w1.transpose(1, 2), w2.transpose(1, 2)
        return w1[:, :inter].contiguous(), w2.contiguous(), w1[:, inter:].contiguous()
    w3 = _param_of(experts_source, spec.expert_w3_name).detach()[lo:lo + nl]
    if tuple(w1.shape[1:]) != (inter, spec.hidden_size):
        w1, w2, w3 = w1.transpose(1, 2), w2.transpose(1, 2), w3.transpose(1, 2)
    return w1.contiguous(), w2.contiguous(), w3.contiguous()


def apply_autoep(model, ep_config, tp_group=None):
    """Replace every detected MoE block with :class:`AutoEPMoELayer` (``expert_parallel`` ds_config section)."""
    from . import auto_ep_presets as P
    cfg = ep_config if isinstance(ep_config, P.AutoEPConfig) else P.parse_autoep_config(ep_config)
    cfg.enabled = True
    P.fill_autoep_config_from_hf(cfg, getattr(model, "config", None))
    auto = AutoEP(model, cfg)
    auto.replace_moe_layers(auto.ep_parser(), tp_group=tp_group)
    return model
