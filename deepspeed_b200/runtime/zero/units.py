"""Flat units: the single sharding primitive shared by ZeRO stages 0-3.

A *unit* is an ordered set of parameters flattened into one padded buffer that is split evenly
("rank-interleaved") over a shard group: rank ``r`` owns ``[r*S, (r+1)*S)`` of every unit. One unit is
exactly one all-gather (params) and one reduce-scatter (grads), so the same object serves as

* the IPG/reduce bucket of ZeRO-1/2 (reference ``stage_1_and_2.py:1193-1468``),
* the per-submodule fetch/release granule of ZeRO-3 (reference ``partitioned_param_coordinator.py:310-490``;
  here the granule is a transformer block, not every ``nn.Module``),
* the sub-group of the partitioned optimizer step (reference ``stage3.py:2629``).

Layout rules that matter to the sm_100a kernels: every parameter starts on a 128-byte boundary and every
shard is a multiple of 256 bytes, so TMA / 16-byte vector accesses never straddle a parameter edge.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

PARAM_ALIGN = 64  # elements: 128 B for 16-bit params
SHARD_ALIGN = 128  # elements per rank shard granularity


def _align(x, a):
    return (x + a - 1) // a * a


@dataclass
class GroupRun:
    """A contiguous range of a unit that belongs to one optimizer param group."""
    group_idx: int
    start: int
    end: int


class FlatUnit:

    def __init__(self, uid, params, names, param_group_ids, world, rank, module=None, persistent=False,
                 dp_group=None, group_key="dp", name=None):
        self.uid = uid
        self.name = name or f"unit{uid}"
        self.module = module
        self.persistent = persistent
        self.dp_group = dp_group
        self.group_key = group_key
        self.world = max(1, world)
        self.rank = rank
        # keep optimizer-group runs contiguous inside the unit
        order = sorted(range(len(params)), key=lambda i: (param_group_ids[i], i))
        self.params = [params[i] for i in order]
        self.names = [names[i] for i in order]
        self.param_group_ids = [param_group_ids[i] for i in order]
        self.shapes = [tuple(p.ds_shape) if hasattr(p, "ds_shape") else tuple(p.shape) for p in self.params]
        self.numels = [int(torch.Size(s).numel()) for s in self.shapes]
        self.dtype = self.params[0].dtype
        self.device = self.params[0].device
        self.offsets = []
        off = 0
        for n in self.numels:
            self.offsets.append(off)
            off = _align(off + n, PARAM_ALIGN)
        self.numel = off
        self.numel_padded = _align(max(off, 1), self.world * SHARD_ALIGN)
        self.shard_numel = self.numel_padded // self.world
        self.runs: List[GroupRun] = []
        for i, g in enumerate(self.param_group_ids):
            end = self.offsets[i + 1] if i + 1 < len(self.params) else self.numel_padded
            if self.runs and self.runs[-1].group_idx == g:
                self.runs[-1].end = end
            else:
                self.runs.append(GroupRun(g, self.offsets[i], end))
        # runtime state
        self.full_param: Optional[torch.Tensor] = None  # flat low-precision params (all ranks' shards)
        self.full_grad: Optional[torch.Tensor] = None  # flat grads prior to reduction
        self.param_shard: Optional[torch.Tensor] = None  # lp shard owned by this rank (view into flat storage)
        self.master_shard: Optional[torch.Tensor] = None  # fp32 shard
        self.grad_shard: Optional[torch.Tensor] = None  # reduced grad shard (accumulation dtype)
        self.shard_offset = 0  # offset of this unit's shard in the rank-flat storage
        self.gathered = False
        self.gather_work = None
        self.gather_event = None
        self.grads_ready = 0
        self.grads_expected = sum(1 for p in self.params if p.requires_grad)
        self.reduced_this_pass = False
        self.has_grad_data = False
        self.sec_shard = None  # hpZ secondary (intra-group) replica shard
        self.opt_event = None  # pipelined optimizer step: recorded after this unit's update on the side stream
        self._ready_epoch = 0  # ... and the symmetric-memory epoch at which every peer has published its shard

    # ---- geometry helpers --------------------------------------------------------------------
    def shard_range(self, rank=None):
        r = self.rank if rank is None else rank
        return r * self.shard_numel, (r + 1) * self.shard_numel

    def runs_in_shard(self, rank=None):
        """Group runs clipped to this rank's shard, in shard-local coordinates."""
        s, e = self.shard_range(rank)
        out = []
        for run in self.runs:
            a, b = max(run.start, s), min(run.end, e)
            if a < b:
                out.append(GroupRun(run.group_idx, a - s, b - s))
        return out

    def param_fragments(self, rank=None):
        """For each param: (param_index, start_in_param, start_in_shard, numel) of the overlap with the shard."""
        s, e = self.shard_range(rank)
        out = []
        for i, (off, n) in enumerate(zip(self.offsets, self.numels)):
            a, b = max(off, s), min(off + n, e)
            if a < b:
                out.append((i, a - off, a - s, b - a))
        return out

    def view_params(self, flat):
        """Per-param views (original shapes) into a full flat buffer."""
        return [flat[o:o + n].view(s) for o, n, s in zip(self.offsets, self.numels, self.shapes)]

    def __repr__(self):
        return (f"FlatUnit({self.name}, params={len(self.params)}, numel={self.numel}, padded={self.numel_padded}, "
                f"shard={self.shard_numel}, persistent={self.persistent}, group={self.group_key})")


def _param_group_index(optimizer_param_groups):
    idx = {}
    for gi, g in enumerate(optimizer_param_groups):
        for p in g["params"]:
            idx[id(p)] = gi
    return idx


def _group_key_of(p):
    """Expert params are sharded/reduced over their expert-data-parallel group (reference moe/utils.py:72)."""
    if hasattr(p, "allreduce") and not p.allreduce and hasattr(p, "group_name"):
        return p.group_name
    return "dp"


def build_bucket_units(named_params, param_groups, bucket_numel, world_of, rank_of, group_of):
    """ZeRO-0/1/2 policy: pack params (reverse order ~ backward order) into buckets <= bucket_numel;
    buckets never mix dtypes or communication groups."""
    pg_idx = _param_group_index(param_groups)
    units, cur, cur_key, cur_numel = [], [], None, 0

    def flush():
        nonlocal cur, cur_numel
        if cur:
            key = cur_key[0]
            units.append(
                FlatUnit(len(units), [p for _, p in cur], [n for n, _ in cur], [pg_idx[id(p)] for _, p in cur],
                         world_of(key), rank_of(key), dp_group=group_of(key), group_key=key, persistent=True))
        cur, cur_numel = [], 0

    for name, p in reversed(list(named_params)):
        if id(p) not in pg_idx:
            continue
        key = (_group_key_of(p), p.dtype, p.device)
        n = p.numel()
        if cur and (key != cur_key or cur_numel + n > bucket_numel):
            flush()
        cur_key = key
        cur.append((name, p))
        cur_numel += n
    flush()
    return units


def build_module_units(model, param_groups, unit_max_numel, persistence_threshold, world_of, rank_of, group_of,
                       leaf_modules=()):
    """ZeRO-3 policy: a module becomes a unit when all its (unassigned) params fit in ``unit_max_numel``;
    otherwise recurse and give its direct params their own unit. Params <= persistence threshold are kept
    replicated in persistent units (reference ``parameter_offload.py:311 mark_persistent_parameters``)."""
    pg_idx = _param_group_index(param_groups)
    names = {id(p): n for n, p in model.named_parameters()}
    assigned = set()
    specs = []  # (module, [params])
    persistent = []

    def numel_of(p):
        return int(torch.Size(p.ds_shape).numel()) if hasattr(p, "ds_shape") else p.numel()

    def is_persistent(p):
        return numel_of(p) <= persistence_threshold

    def unassigned(module, recurse=True):
        out, seen = [], set()
        for p in module.parameters(recurse=recurse):
            if id(p) in assigned or id(p) in seen or id(p) not in pg_idx:
                continue
            seen.add(id(p))
            out.append(p)
        return out

    def take(module, ps):
        big = []
        for p in ps:
            assigned.add(id(p))
            (persistent if is_persistent(p) else big).append(p)
        # split by (comm group, dtype)
        by_key = {}
        for p in big:
            by_key.setdefault((_group_key_of(p), p.dtype), []).append(p)
        for key, plist in by_key.items():
            specs.append((module, plist, key[0]))

    def visit(module, is_root):
        ps = unassigned(module)
        if not ps:
            return
        total = sum(numel_of(p) for p in ps)
        children = list(module.children())
        leaf = bool(getattr(module, "_z3_leaf", False)) or (isinstance(module, tuple(leaf_modules))
                                                           if leaf_modules else False)
        if leaf:
            for p in module.parameters():
                p.ds_z3_leaf_module = module
        container = isinstance(module, (torch.nn.ModuleList, torch.nn.ModuleDict))  # no forward(): hooks never fire
        if container:
            for c in children:
                visit(c, False)
            return
        if leaf or not children or (total <= unit_max_numel and not is_root) or (is_root and total <= min(
                unit_max_numel, persistence_threshold)):
            take(module, ps)
            return
        for c in children:
            visit(c, False)
        direct = unassigned(module, recurse=False)
        if direct:
            take(module, direct)

    visit(model, True)
    units = []
    for module, plist, key in specs:
        units.append(
            FlatUnit(len(units), plist, [names[id(p)] for p in plist], [pg_idx[id(p)] for p in plist], world_of(key),
                     rank_of(key), module=module, dp_group=group_of(key), group_key=key, persistent=False))
    by_key = {}
    for p in persistent:
        by_key.setdefault((_group_key_of(p), p.dtype), []).append(p)
    for key, plist in by_key.items():
        units.append(
            FlatUnit(len(units), plist, [names[id(p)] for p in plist], [pg_idx[id(p)] for p in plist],
                     world_of(key[0]), rank_of(key[0]), module=None, dp_group=group_of(key[0]), group_key=key[0],
                     persistent=True, name=f"persistent{len(units)}"))
    return units
