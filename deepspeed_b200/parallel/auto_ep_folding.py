"""AutoEP folded over AutoTP: expert-parallel groups laid across tensor-parallel lanes.

Reference: ``module_inject/auto_ep_folding.py`` (``ParallelFoldingSpec``, ``expected_folding_group_tables``,
``validate_folding_global``, the gradient-family rules) and ``moe/ep_tp_dispatch.py`` (token partition across the TP
peers before the expert dispatch, restore after the combine).

With tensor parallelism the tokens entering an MoE block are replicated on the ``tp`` lanes of a replica. Unfolded,
every lane would route and dispatch the same tokens (``tp`` times the expert work); folded, lane ``l`` routes only
rows ``[l*T/tp, (l+1)*T/tp)``, the expert-parallel group spans ``ep`` consecutive ranks of the stage (so it may cross
lanes), and the outputs are all-gathered over the lanes afterwards.

Gradient design (differs from the reference, which all-gathers with a summing backward and then divides gradients by
``tp`` per parameter family): the two autograd functions here are exact adjoints of the *replicated* computation -
``partition_tokens`` slices in forward and all-gathers in backward, ``restore_tokens`` all-gathers in forward and
slices in backward - so activations' gradients come out replicated and unscaled, routed-expert gradients need nothing
beyond their expert-data-parallel reduction, and the only partial gradients are those of the router (each lane saw a
slice of the tokens): ``sum_grad_over_lanes`` wraps the router weight so that its gradient is summed over the lanes
inside autograd. No post-hoc correction pass over the gradient buffers is needed.
"""
from dataclasses import dataclass
from typing import Dict, Tuple

import torch

from .. import comm as dist

GRAD_REDUCE_SKIP = "skip"
GRAD_REDUCE_SUM = "sum"
GRAD_REDUCE_AVERAGE = "average"
PARAM_FAMILY_ATTR = "ds_autoep_folding_param_family"
ROUTER_GATE_PARTIAL = "router_gate_partial"


@dataclass(frozen=True)
class ParallelFoldingSpec:
    world_size: int
    pp_size: int
    stage_size: int
    tp_size: int
    dp_size: int
    ep_size: int
    etp_size: int
    edp_size: int
    mp_mode: str = "tp"


@dataclass(frozen=True)
class FoldingGroupTables:
    tp_groups: Tuple[Tuple[int, ...], ...]
    dense_dp_groups: Tuple[Tuple[int, ...], ...]
    ep_groups: Tuple[Tuple[int, ...], ...]
    edp_groups: Tuple[Tuple[int, ...], ...]


@dataclass(frozen=True)
class FoldingGroupHandles:
    spec: ParallelFoldingSpec
    tp_group: object
    dense_dp_group: object
    ep_group: object
    edp_group: object
    ep_group_name: str
    tp_ranks: Tuple[int, ...]
    dense_dp_ranks: Tuple[int, ...]
    ep_ranks: Tuple[int, ...]
    edp_ranks: Tuple[int, ...]


def _divisors(n):
    return [d for d in range(1, n + 1) if n % d == 0]


def build_folding_spec(*, world_size, pp_size=1, tp_size=1, ep_size=1, etp_size=1, mp_mode="tp"):
    for name, v in (("world_size", world_size), ("pp_size", pp_size), ("tensor_parallel.autotp_size", tp_size),
                    ("expert_parallel.autoep_size", ep_size),
                    ("expert_parallel.expert_tensor_parallel_size", etp_size)):
        if not isinstance(v, int) or isinstance(v, bool) or v < 1:
            raise ValueError(f"{name} must be a positive integer, got {v!r}")
    if world_size % pp_size:
        raise ValueError(f"pp_size={pp_size} must divide world_size={world_size}. Valid pp_size values: "
                         f"{_divisors(world_size)}")
    stage = world_size // pp_size
    if stage % tp_size:
        raise ValueError(f"tensor_parallel.autotp_size={tp_size} must divide the stage size ({stage}). Valid "
                         f"autotp_size values: {_divisors(stage)}")
    width = ep_size * etp_size
    if stage % width:
        raise ValueError(f"expert_parallel.autoep_size * expert_parallel.expert_tensor_parallel_size ({ep_size} * "
                         f"{etp_size} = {width}) must divide the stage size ({stage}). Valid expert-width values: "
                         f"{_divisors(stage)}")
    return ParallelFoldingSpec(world_size, pp_size, stage, tp_size, stage // tp_size, ep_size, etp_size,
                               stage // width, mp_mode)


def expected_folding_group_tables(spec) -> FoldingGroupTables:
    """Rank tables (no process groups): TP lanes are consecutive ranks, dense DP is strided by ``tp``; EP groups are
    ``ep`` consecutive ranks of a stage, expert-DP groups take the same position of every EP group."""
    tp, ddp, ep, edp = [], [], [], []
    for s0 in range(0, spec.world_size, spec.stage_size):
        ranks = list(range(s0, s0 + spec.stage_size))
        tp += [tuple(ranks[i * spec.tp_size:(i + 1) * spec.tp_size]) for i in range(spec.dp_size)]
        ddp += [tuple(ranks[l::spec.tp_size]) for l in range(spec.tp_size)]
        local = [tuple(ranks[i:i + spec.ep_size]) for i in range(0, len(ranks), spec.ep_size)]
        ep += local
        edp += [tuple(g[p] for g in local) for p in range(spec.ep_size)]
    return FoldingGroupTables(tuple(tp), tuple(ddp), tuple(ep), tuple(edp))


def local_folding_ranks(global_rank, spec) -> Dict[str, Tuple[int, ...]]:
    t = expected_folding_group_tables(spec)
    return {k: next(g for g in gs if global_rank in g)
            for k, gs in (("tp", t.tp_groups), ("dense_dp", t.dense_dp_groups), ("ep", t.ep_groups),
                          ("edp", t.edp_groups))}


def _mpu_size(mpu, *names):
    for n in names:
        f = getattr(mpu, n, None) if mpu is not None else None
        if f is not None:
            return f()
    return None


def validate_folding_global(spec, *, zero_stage=0, sp_size=1, deepcompile_enabled=False,
                            use_data_before_expert_parallel=False, mpu=None, autoep_enabled=True, tp_preset=None,
                            ep_preset=None, zero_offload_optimizer=False, zero_offload_param=False):
    """Policy checks before any process group exists."""
    if not autoep_enabled:
        return
    folded = spec.tp_size > 1
    if spec.etp_size != 1:
        raise ValueError(f"expert_parallel.expert_tensor_parallel_size={spec.etp_size}: tensor parallelism inside an "
                         "expert is not supported, use 1")
    if tp_preset is not None and ep_preset is not None and tp_preset != ep_preset:
        raise ValueError(f"tensor_parallel.preset_model ({tp_preset!r}) and expert_parallel.preset_model "
                         f"({ep_preset!r}) must match when both are set")
    if folded:
        if deepcompile_enabled:
            raise ValueError("DeepCompile with AutoEP+AutoTP folding is not supported")
        if spec.pp_size > 1:
            raise ValueError(f"AutoEP+AutoTP folding supports pp_size=1 only, got pp_size={spec.pp_size}")
        if sp_size > 1:
            raise ValueError(f"tensor parallelism (autotp_size={spec.tp_size}) and Ulysses sequence parallelism "
                             f"(sp_size={sp_size}) are mutually exclusive under AutoEP folding")
        if spec.ep_size == 1:
            raise ValueError("AutoEP+AutoTP folding requires expert_parallel.autoep_size > 1")
        if use_data_before_expert_parallel:
            raise ValueError("use_data_before_expert_parallel_ is not supported with AutoEP+AutoTP folding")
        if zero_stage == 3:
            raise ValueError("AutoEP+AutoTP folding supports ZeRO stage 0, 1 or 2")
        if zero_offload_optimizer or zero_offload_param:
            raise ValueError("ZeRO offload with AutoEP+AutoTP folding is not supported")
    mpu_tp = _mpu_size(mpu, "get_tensor_model_parallel_world_size", "get_model_parallel_world_size")
    if mpu_tp not in (None, 1, spec.tp_size):
        raise ValueError(f"mpu tensor/model parallel world size ({mpu_tp}) conflicts with autotp_size={spec.tp_size}")
    mpu_pp = _mpu_size(mpu, "get_pipeline_model_parallel_world_size", "get_pipeline_parallel_world_size")
    if mpu_pp not in (None, spec.pp_size):
        raise ValueError(f"mpu pipeline parallel world size ({mpu_pp}) conflicts with pp_size={spec.pp_size}")


def assert_group_matches_spec(existing_rank_lists, spec):
    """Cached ``ep_size_N`` rank lists (``(ep_lists, edp_lists)`` or a dict) must be the ones the spec asks for."""
    t = expected_folding_group_tables(spec)
    if isinstance(existing_rank_lists, dict):
        got_ep, got_edp = existing_rank_lists.get("ep", []), existing_rank_lists.get("edp", [])
    else:
        got_ep, got_edp = existing_rank_lists
    for kind, got, want in (("expert-parallel", got_ep, set(t.ep_groups)),
                            ("expert-data-parallel", got_edp, set(t.edp_groups))):
        for g in got:
            if tuple(int(r) for r in g) not in want:
                raise RuntimeError(f"Cached {kind} group {tuple(g)} does not match folding spec {spec}")


def create_folding_groups(spec) -> FoldingGroupHandles:
    """Build the four process groups of this rank (collective: every rank creates every group)."""
    t = expected_folding_group_tables(spec)
    rank = dist.get_rank()

    def make(rank_lists):
        mine, mine_ranks = None, None
        for rl in rank_lists:
            g = dist.new_group(list(rl)) if len(rl) > 1 else None
            if rank in rl:
                mine, mine_ranks = g, tuple(rl)
        return mine, mine_ranks

    tp_g, tp_r = make(t.tp_groups)
    ddp_g, ddp_r = make(t.dense_dp_groups)
    ep_g, ep_r = make(t.ep_groups)
    edp_g, edp_r = make(t.edp_groups)
    return FoldingGroupHandles(spec, tp_g, ddp_g, ep_g, edp_g, f"ep_size_{spec.ep_size}", tp_r, ddp_r, ep_r, edp_r)


# ---- token partition over the TP lanes ----------------------------------------------------------------------
def _rows_per_lane(total, lanes):
    return (total + lanes - 1) // lanes


class _PartitionTokens(torch.autograd.Function):
    """Forward: this lane's row block (zero rows appended when ``T`` is not a multiple of the lanes). Backward: the
    lanes' row-gradients gathered back into the replicated ``[T, H]`` gradient."""

    @staticmethod
    def forward(ctx, x, group, lane, lanes):
        ctx.group, ctx.lane, ctx.lanes, ctx.total = group, lane, lanes, x.shape[0]
        per = _rows_per_lane(x.shape[0], lanes)
        lo, hi = min(lane * per, x.shape[0]), min((lane + 1) * per, x.shape[0])
        out = x.new_zeros((per, ) + tuple(x.shape[1:]))
        out[:hi - lo] = x[lo:hi]
        return out

    @staticmethod
    def backward(ctx, g):
        full = g.new_empty((ctx.lanes * g.shape[0], ) + tuple(g.shape[1:]))
        dist.all_gather_into_tensor(full, g.contiguous(), group=ctx.group)
        return full[:ctx.total], None, None, None


class _RestoreTokens(torch.autograd.Function):
    """Forward: all-gather the lanes' row blocks into ``[T, H]``. Backward: this lane's rows of the (replicated)
    gradient - no reduction, the other lanes hold the same gradient."""

    @staticmethod
    def forward(ctx, y, group, lane, lanes, total):
        ctx.lane, ctx.per, ctx.total = lane, y.shape[0], total
        full = y.new_empty((lanes * y.shape[0], ) + tuple(y.shape[1:]))
        dist.all_gather_into_tensor(full, y.contiguous(), group=group)
        return full[:total]

    @staticmethod
    def backward(ctx, g):
        lo, hi = min(ctx.lane * ctx.per, ctx.total), min((ctx.lane + 1) * ctx.per, ctx.total)
        out = g.new_zeros((ctx.per, ) + tuple(g.shape[1:]))
        out[:hi - lo] = g[lo:hi]
        return out, None, None, None, None


class _SumGradOverLanes(torch.autograd.Function):

    @staticmethod
    def forward(ctx, w, group):
        ctx.group = group
        return w.view_as(w)

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().clone()
        dist.all_reduce(g, group=ctx.group)
        return g, None


def partition_tokens(x, tp_group, lane=None, lanes=None):
    """``[T, H]`` replicated on the lanes -> ``[ceil(T/tp), H]`` of this lane, and the number of real rows in it."""
    lanes = lanes if lanes is not None else dist.get_world_size(group=tp_group)
    lane = lane if lane is not None else dist.get_rank(group=tp_group)
    per = _rows_per_lane(x.shape[0], lanes)
    real = max(0, min((lane + 1) * per, x.shape[0]) - lane * per)
    return _PartitionTokens.apply(x, tp_group, lane, lanes), real


def restore_tokens(y, total, tp_group, lane=None, lanes=None):
    lanes = lanes if lanes is not None else dist.get_world_size(group=tp_group)
    lane = lane if lane is not None else dist.get_rank(group=tp_group)
    return _RestoreTokens.apply(y, tp_group, lane, lanes, total)


def sum_grad_over_lanes(weight, tp_group):
    """Identity whose gradient is all-reduced (sum) over the lanes: for parameters that saw only a slice of the
    replicated tokens (the folded router)."""
    return _SumGradOverLanes.apply(weight, tp_group)


def mark_autoep_folding_partial_router_parameter(param):
    setattr(param, PARAM_FAMILY_ATTR, ROUTER_GATE_PARTIAL)


def autoep_folding_gradient_reduction_strategy(folding_spec, param, *, param_name=None, reduced_in_autograd=True):
    """What an engine still has to do over the TP lanes for ``param``: nothing for TP-sharded and routed-expert
    parameters and for everything replicated (identical gradients on every lane); ``sum`` for a router that saw a
    token slice - unless its weight went through :func:`sum_grad_over_lanes` (``reduced_in_autograd``), which is what
    :class:`AutoEPMoELayer` does."""
    if folding_spec is None or getattr(folding_spec, "tp_size", 1) <= 1:
        return GRAD_REDUCE_SKIP
    if getattr(param, "model_parallel", False) or getattr(param, "tensor_model_parallel", False):
        return GRAD_REDUCE_SKIP
    if hasattr(param, "allreduce") and not param.allreduce:
        return GRAD_REDUCE_SKIP
    if getattr(param, PARAM_FAMILY_ATTR, None) == ROUTER_GATE_PARTIAL and not reduced_in_autograd:
        return GRAD_REDUCE_SUM
    return GRAD_REDUCE_SKIP
