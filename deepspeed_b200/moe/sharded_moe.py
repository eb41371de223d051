"""GShard-style capacity MoE: TopKGate (top-1/2/k), MOELayer (dispatch -> experts -> combine).

Reference behaviour: ``deepspeed/moe/sharded_moe.py`` (``top1gating`` :235, ``top2gating`` :341, ``topkgating``
:434, ``TopKGate`` :528, ``MOELayer`` :618-774, ``_AllToAll`` :104). Capacity = ceil(tokens/experts * cf)
(min ``min_capacity``); aux load-balancing loss ``sum(me*ce)*E``; RTS (random token selection) and jitter noise;
``drop_tokens=False`` syncs capacity with an all-reduce MAX.

Re-design: instead of one-hot einsum dispatch (``[S,E,C]`` masks -> ``einsum("sec,sm->ecm")``) tokens are
*index-routed*: a slot index per (token, k) is computed and tokens are gathered/scattered with index ops, which
is O(S*M) instead of O(S*E*C*M). The all-to-all is ``deepspeed_b200.comm.all_to_all_single`` on the baseline
path and the fused NVLink dispatch/combine kernel (:mod:`deepspeed_b200.comm.symm`) when available.
"""
import math
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from .. import comm as dist

uniform_map = {}
gumbel_map = {}
exp_selection_uniform_map = {}


def multiplicative_jitter(x, device, epsilon=1e-2):
    if epsilon == 0:
        return x
    uniform = uniform_map.get(device)
    if uniform is None:
        uniform = torch.distributions.uniform.Uniform(low=torch.tensor(1.0 - epsilon, device=device),
                                                      high=torch.tensor(1.0 + epsilon, device=device)).rsample
        uniform_map[device] = uniform
    return x * uniform(x.shape)


def gumbel_rsample(shape, device):
    gumbel = gumbel_map.get(device)
    if gumbel is None:
        one, zero = torch.tensor(1.0, device=device), torch.tensor(0.0, device=device)
        gumbel = torch.distributions.gumbel.Gumbel(zero, one).rsample
        gumbel_map[device] = gumbel
    return gumbel(shape)


class _AllToAll(torch.autograd.Function):

    @staticmethod
    def forward(ctx, group, input):
        ctx.group = group
        input = input.contiguous()
        output = torch.empty_like(input)
        dist.all_to_all_single(output, input, group=group)
        return output

    @staticmethod
    def backward(ctx, grad_output):
        return None, _AllToAll.apply(ctx.group, grad_output)


def _capacity(num_tokens, num_experts, capacity_factor, min_capacity):
    cap = int(math.ceil((num_tokens / num_experts) * capacity_factor))
    return max(cap, int(min_capacity))


def _slots_from_mask(mask):
    """position of each selected token inside its expert queue: cumsum(mask, 0) - 1 (per expert)."""
    return torch.cumsum(mask, dim=0) - 1


def topkgating(logits, k, capacity_factor, min_capacity, drop_tokens=True, ep_group=None, drop_policy="probs",
               noisy_gate_policy=None, use_rts=True, used_token=None, normalize=True):
    """Generic top-k gating.

    Returns ``(l_aux, expert_idx [S,k], slot_idx [S,k] (-1 = dropped), gate_w [S,k], capacity, exp_counts [E])``.
    """
    S, E = logits.shape
    gates = F.softmax(logits.float(), dim=1)
    sel_logits = logits
    if noisy_gate_policy == "RSample":
        sel_logits = logits + gumbel_rsample(logits.shape, device=logits.device)
    top_vals, top_idx = torch.topk(sel_logits if noisy_gate_policy == "RSample" else gates, k=k, dim=1)
    top_gate = gates.gather(1, top_idx)  # [S,k]
    mask = torch.zeros(S, E, dtype=torch.int64, device=logits.device)
    mask.scatter_(1, top_idx, 1)
    if used_token is not None:
        mask = mask * used_token.view(-1, 1).to(mask.dtype)
    exp_counts = mask.sum(dim=0).detach()
    # load-balancing loss (top-k generalisation: divide ce by k)
    me = gates.mean(dim=0)
    ce = mask.float().mean(dim=0)
    l_aux = torch.sum(me * ce) * E / k

    if drop_tokens:
        capacity = _capacity(S * k, E, capacity_factor, min_capacity)
    else:
        new_cap = exp_counts.max()
        if ep_group is not None and dist.is_initialized() and dist.get_world_size(ep_group) > 1:
            dist.all_reduce(new_cap, op=dist.ReduceOp.MAX, group=ep_group)
        capacity = int(new_cap)  # one host sync, same as the reference (sharded_moe.py:686)

    # priority inside each expert queue
    if drop_tokens and use_rts and k == 1:
        uniform = exp_selection_uniform_map.get(logits.device)
        if uniform is None:
            uniform = torch.distributions.uniform.Uniform(low=torch.tensor(0.0, device=logits.device),
                                                          high=torch.tensor(1.0, device=logits.device)).rsample
            exp_selection_uniform_map[logits.device] = uniform
        prio = mask * uniform(mask.shape)
        # keep the `capacity` highest-priority tokens per expert
        keep_idx = torch.topk(prio, k=min(capacity, S), dim=0)[1]
        keep = torch.zeros_like(mask).scatter_(0, keep_idx, 1) * mask
        slots = _slots_from_mask(keep)
        valid = keep.bool()
    elif drop_tokens and drop_policy == "probs" and k > 1:
        prio = mask * gates
        keep_idx = torch.topk(prio, k=min(capacity, S), dim=0)[1]
        keep = torch.zeros_like(mask).scatter_(0, keep_idx, 1) * mask
        slots = _slots_from_mask(keep)
        valid = keep.bool()
    else:  # position policy: earlier tokens win
        slots = _slots_from_mask(mask)
        valid = mask.bool() & (slots < capacity)
    slot_idx = slots.gather(1, top_idx)
    ok = valid.gather(1, top_idx)
    slot_idx = torch.where(ok, slot_idx, torch.full_like(slot_idx, -1))
    w = top_gate * ok.to(top_gate.dtype)
    if normalize and k > 1:
        denom = w.sum(dim=1, keepdim=True).clamp(min=torch.finfo(w.dtype).eps)
        w = w / denom
    return l_aux, top_idx, slot_idx, w, capacity, exp_counts


def top1gating(logits, capacity_factor, min_capacity, used_token=None, noisy_gate_policy=None, drop_tokens=True,
               use_rts=True, ep_group=None, use_tutel=False):
    return topkgating(logits, 1, capacity_factor, min_capacity, drop_tokens, ep_group, "position", noisy_gate_policy,
                      use_rts, used_token, normalize=False)


def top2gating(logits, capacity_factor, min_capacity, drop_tokens=True, ep_group=None, top2_2nd_expert_sampling=True):
    """Top-2 gating with the reference's exact selection rules (``moe/sharded_moe.py:341-404``): expert 1 is the
    noise-free argmax of the gates; expert 2 is the argmax of the (optionally Gumbel-perturbed) logits with expert 1
    masked out; all second choices queue behind all first choices of the same expert
    (``locations2 += sum(mask1)``); the balancing loss uses the first-choice mask only. Same return tuple as
    :func:`topkgating`."""
    S, E = logits.shape
    logits = logits.float()
    gates = F.softmax(logits, dim=1)
    idx1 = torch.argmax(gates, dim=1)
    mask1 = F.one_hot(idx1, num_classes=E)
    sel = logits + gumbel_rsample(logits.shape, device=logits.device) if top2_2nd_expert_sampling else logits
    idx2 = torch.argmax(sel.masked_fill(mask1.bool(), float("-inf")), dim=1)
    mask2 = F.one_hot(idx2, num_classes=E)
    loc1 = torch.cumsum(mask1, dim=0) - 1
    loc2 = torch.cumsum(mask2, dim=0) - 1 + mask1.sum(dim=0, keepdim=True)
    me = gates.mean(dim=0)
    ce = mask1.float().mean(dim=0)
    l_aux = torch.mean(me * ce) * E * E
    exp_counts = (mask1 + mask2).sum(dim=0).detach()
    if drop_tokens:
        capacity = _capacity(S * 2, E, capacity_factor, min_capacity)
    else:
        new_cap = exp_counts.max()
        if ep_group is not None and dist.is_initialized() and dist.get_world_size(ep_group) > 1:
            dist.all_reduce(new_cap, op=dist.ReduceOp.MAX, group=ep_group)
        capacity = int(new_cap)
    slot1 = loc1.gather(1, idx1.view(-1, 1)).squeeze(1)
    slot2 = loc2.gather(1, idx2.view(-1, 1)).squeeze(1)
    ok1, ok2 = slot1 < capacity, slot2 < capacity
    g1 = gates.gather(1, idx1.view(-1, 1)).squeeze(1) * ok1.to(gates.dtype)
    g2 = gates.gather(1, idx2.view(-1, 1)).squeeze(1) * ok2.to(gates.dtype)
    denom = (g1 + g2).clamp(min=torch.finfo(gates.dtype).eps)
    top_idx = torch.stack([idx1, idx2], dim=1)
    slot_idx = torch.stack([torch.where(ok1, slot1, torch.full_like(slot1, -1)),
                            torch.where(ok2, slot2, torch.full_like(slot2, -1))], dim=1)
    w = torch.stack([g1 / denom, g2 / denom], dim=1)
    return l_aux, top_idx, slot_idx, w, capacity, exp_counts


class TopKGate(nn.Module):
    """Gate network: ``wg`` linear (fp32) + top-k gating."""

    def __init__(self, model_dim, num_experts, k=1, capacity_factor=1.0, eval_capacity_factor=1.0, min_capacity=8,
                 noisy_gate_policy: Optional[str] = None, drop_tokens=True, use_rts=True, ep_group=None,
                 top2_2nd_expert_sampling=True):
        super().__init__()
        self.wg = nn.Linear(model_dim, num_experts, bias=False)
        self.ep_group = ep_group
        self.k = k
        self.capacity_factor = capacity_factor
        self.eval_capacity_factor = eval_capacity_factor
        self.min_capacity = min_capacity
        self.noisy_gate_policy = noisy_gate_policy
        self.wall_clock_breakdown = False
        self.gate_time = 0.0
        self.drop_tokens = drop_tokens
        self.use_rts = use_rts
        self.top2_2nd_expert_sampling = top2_2nd_expert_sampling

    def _set_ep_group(self, ep_group):
        assert self.ep_group is None, "Attempting to override an existing ep_group"
        self.ep_group = ep_group

    def forward(self, input, used_token=None, use_tutel=False):
        input_fp32 = input.float()
        if self.noisy_gate_policy == "Jitter" and self.training:
            input_fp32 = multiplicative_jitter(input_fp32, device=input.device)
        logits = F.linear(input_fp32, self.wg.weight.float())
        cf = self.capacity_factor if self.training else self.eval_capacity_factor
        if self.k == 1:
            return top1gating(logits, cf, self.min_capacity, used_token,
                              self.noisy_gate_policy if self.training else None, self.drop_tokens, self.use_rts,
                              self.ep_group)
        if self.k == 2:
            return top2gating(logits, cf, self.min_capacity, self.drop_tokens, self.ep_group,
                              self.top2_2nd_expert_sampling)
        return topkgating(logits, self.k, cf, self.min_capacity, self.drop_tokens, self.ep_group)


class MOELayer(nn.Module):
    """dispatch (index scatter -> all-to-all) -> local experts -> combine (all-to-all -> weighted gather)."""

    def __init__(self, gate, experts, ep_group_name, ep_size, num_local_experts, use_tutel=False):
        super().__init__()
        self.gate = gate
        self.experts = experts
        self.ep_group = None
        self.ep_size = ep_size
        self.ep_group_name = ep_group_name
        self.num_local_experts = num_local_experts
        self.time_falltoall = self.time_salltoall = self.time_moe = 0.0
        self.wall_clock_breakdown = False
        self.l_aux = None
        self.exp_counts = None

    def _set_ep_group(self, ep_group):
        self.ep_group = ep_group
        self.gate._set_ep_group(ep_group)

    def forward(self, *input, **kwargs):
        x = input[0]
        d_model = x.shape[-1]
        tokens = x.reshape(-1, d_model)
        S = tokens.shape[0]
        l_aux, expert_idx, slot_idx, gate_w, capacity, exp_counts = self.gate(tokens, input[1] if len(input) > 1 else None)
        self.l_aux, self.exp_counts = l_aux, exp_counts
        E = self.ep_size * self.num_local_experts
        k = expert_idx.shape[1]
        # ---- dispatch: [E*C, M] buffer, token (s,j) -> row expert*C + slot
        valid = slot_idx >= 0
        rows = (expert_idx * capacity + slot_idx.clamp(min=0)).reshape(-1)
        src_tok = torch.arange(S, device=x.device).repeat_interleave(k)
        vflat = valid.reshape(-1)
        rows_v, src_v = rows[vflat], src_tok[vflat]
        dispatched = torch.zeros(E * capacity, d_model, dtype=tokens.dtype, device=tokens.device)
        dispatched.index_copy_(0, rows_v, tokens.index_select(0, src_v))
        # tensor-parallel ranks hold the same tokens: unless the experts are tensor-parallel themselves, each rank
        # sends only its 1/tp slice of every expert's capacity through the all-to-all (reference sharded_moe.py:586)
        full_capacity = capacity
        tp_sliced = not getattr(self, "enable_expert_tensor_parallelism", False)
        if tp_sliced:
            from .mappings import drop_tokens as _drop, gather_tokens as _gather, _tp
            tp_world = _tp()[0]
            tp_sliced = tp_world > 1 and capacity % tp_world == 0
        if tp_sliced:
            dispatched = _drop(dispatched.view(E, capacity, d_model), dim=1)
            capacity = capacity // tp_world
            dispatched = dispatched.reshape(E * capacity, d_model)
        if self.ep_size > 1:
            dispatched = _AllToAll.apply(self.ep_group, dispatched)
        # [ep, E_local, C, M] -> [E_local, ep*C, M]
        dispatched = dispatched.reshape(self.ep_size, self.num_local_experts, capacity, d_model)
        expert_in = dispatched.permute(1, 0, 2, 3).reshape(self.num_local_experts, self.ep_size * capacity, d_model)
        expert_out = self.experts(expert_in)
        expert_out = expert_out.reshape(self.num_local_experts, self.ep_size, capacity, d_model).permute(1, 0, 2, 3)
        expert_out = expert_out.reshape(E * capacity, d_model)
        if self.ep_size > 1:
            expert_out = _AllToAll.apply(self.ep_group, expert_out)
        if tp_sliced:
            expert_out = _gather(expert_out.view(E, capacity, d_model), dim=1).reshape(E * full_capacity, d_model)
            capacity = full_capacity
        # ---- combine: out[s] = sum_j w[s,j] * expert_out[row(s,j)]
        picked = expert_out.index_select(0, rows_v) * gate_w.reshape(-1)[vflat].unsqueeze(1).to(expert_out.dtype)
        out = torch.zeros(S, d_model, dtype=expert_out.dtype, device=x.device)
        out.index_add_(0, src_v, picked)
        return out.reshape(x.shape)
