"""Translate between the runtime flat-unit shard layout and the reference on-disk ZeRO layouts.

Runtime layout (``runtime/zero/units.py``): every unit is rank-interleaved, 128-byte aligned -- good for TMA and
for one-collective-per-unit, but not what ``zero_to_fp32.py`` / ``ds_to_universal.py`` expect. The reference
formats (``stage_1_and_2.py:2594-2638``, ``stage3.py:3165-3181``, ``utils/zero_to_fp32.py:282-460``):

* stage 1/2: per optimizer group, all params flattened back-to-back, padded to ``2*world``, rank ``r`` holds the
  contiguous slice ``[r*P,(r+1)*P)`` (``single_partition_of_fp32_groups``; last rank's padding stripped), plus
  ``param_slice_mappings`` (name -> (start, numel) inside the rank partition).
* stage 3: per parameter ``ceil(numel/world)`` elements per rank, concatenated in ``param_shapes`` order
  (``fp32_flat_groups``).

Export gathers one unit at a time (peak extra memory = one fp32 unit) and copies the overlaps; import is the
inverse (scatter into a zeroed unit + all-reduce), so only same-rank files are read.
"""
import math
from collections import OrderedDict
from dataclasses import dataclass

import torch

from .. import comm as dist


from ..utils.tensor_fragment import fragment_address  # noqa: E402  (pickled under the reference's module path)


def _cdiv(a, b):
    return (a + b - 1) // b


class ReferenceLayout:
    """Geometry of the reference flat layouts for one :class:`ZeroShardedOptimizer`."""

    def __init__(self, zopt):
        self.z = zopt
        self.world = zopt._shard_world
        self.rank = zopt._shard_rank
        self.stage = zopt.stage
        n_groups = len(zopt.optimizer.param_groups)
        self.group_params = [[] for _ in range(n_groups)]  # [(name, param, numel)]
        for gi, plist in enumerate(zopt._orig_param_groups):
            for p in plist:
                ent = zopt._unit_of_param.get(id(p))
                if ent is None:
                    continue
                u, i = ent
                if getattr(u, "group_key", "dp") != "dp":
                    continue  # expert units: sharded over their own (expert-data-parallel) group, saved per unit
                self.group_params[gi].append((u.names[i], p, u.numels[i], tuple(u.shapes[i])))
        # offsets of every param in the reference flat of its group
        self.ref_offset = {}
        self.group_numel = []
        self.group_padded = []
        self.partition_size = []
        for gi, plist in enumerate(self.group_params):
            off = 0
            for name, p, n, _ in plist:
                if self.stage >= 3:
                    self.ref_offset[id(p)] = (gi, off)  # offset inside the *rank* flat
                    off += _cdiv(n, self.world)
                else:
                    self.ref_offset[id(p)] = (gi, off)
                    off += n
            self.group_numel.append(off)
            if self.stage >= 3:
                self.group_padded.append(off)
                self.partition_size.append(off)
            else:
                align = 2 * self.world
                padded = _cdiv(max(off, 1), align) * align
                self.group_padded.append(padded)
                self.partition_size.append(padded // self.world)

    # ---- metadata ------------------------------------------------------------------------------
    def param_shapes(self):
        return [OrderedDict((name, torch.Size(shape)) for name, _, _, shape in plist) for plist in self.group_params]

    def group_paddings(self):
        out = []
        for gi in range(len(self.group_params)):
            if self.stage >= 3:
                out.append(0)
            else:
                out.append(self.group_padded[gi] - self.group_numel[gi] if self.rank == self.world - 1 else 0)
        return out

    def param_slice_mappings(self):
        maps = []
        for gi, plist in enumerate(self.group_params):
            m = OrderedDict()
            P = self.partition_size[gi]
            lo, hi = self.rank * P, (self.rank + 1) * P
            for name, p, n, _ in plist:
                _, off = self.ref_offset[id(p)]
                a, b = max(off, lo), min(off + n, hi)
                if a < b:
                    m[name] = fragment_address(numel=b - a, start=a - lo)
            maps.append(m)
        return maps

    # ---- export --------------------------------------------------------------------------------
    def _full_unit(self, u, shard):
        if u.world <= 1:
            return shard.to(u.device if shard.device.type != "cpu" or u.device.type == "cpu" else u.device)
        s = shard.to(u.device).contiguous()
        out = torch.empty(u.numel_padded, dtype=s.dtype, device=u.device)
        dist.all_gather_into_tensor(out, s, group=u.dp_group)
        return out

    @torch.no_grad()
    def export(self, shard_of_unit):
        """``shard_of_unit(u) -> fp32 shard tensor``; returns list (per group) of this rank's reference partition."""
        outs = [torch.zeros(self.partition_size[gi], dtype=torch.float32) for gi in range(len(self.group_params))]
        for u in self.z.units:
            if not any(id(p) in self.ref_offset for p in u.params):
                continue  # nothing of this unit is exported (e.g. all parameters frozen); same decision on every rank
            shard = shard_of_unit(u)
            if shard is None:  # this rank's shard holds no such state (only frozen runs here): peers still gather
                shard = torch.zeros(u.shard_numel, dtype=torch.float32, device=u.master_shard.device)
            full = self._full_unit(u, shard.float())
            for i, p in enumerate(u.params):
                if id(p) not in self.ref_offset:
                    continue
                gi, off = self.ref_offset[id(p)]
                n = u.numels[i]
                src = full[u.offsets[i]:u.offsets[i] + n]
                if self.stage >= 3:
                    per = _cdiv(n, self.world)
                    a, b = self.rank * per, min((self.rank + 1) * per, n)
                    if b > a:
                        outs[gi][off:off + (b - a)].copy_(src[a:b])
                else:
                    P = self.partition_size[gi]
                    lo, hi = self.rank * P, (self.rank + 1) * P
                    a, b = max(off, lo), min(off + n, hi)
                    if a < b:
                        outs[gi][a - lo:b - lo].copy_(src[a - off:b - off])
            del full
        return outs

    # ---- import --------------------------------------------------------------------------------
    @torch.no_grad()
    def import_(self, partitions, write_shard):
        """Inverse of :meth:`export`. ``partitions``: list per group of this rank's reference partition;
        ``write_shard(u, fp32_shard_tensor)`` stores the unit shard."""
        for u in self.z.units:
            if not any(id(p) in self.ref_offset for p in u.params):
                continue  # not part of the reference flats (expert units, all-frozen units)
            full = torch.zeros(u.numel_padded, dtype=torch.float32, device=u.device)
            for i, p in enumerate(u.params):
                if id(p) not in self.ref_offset:
                    continue
                gi, off = self.ref_offset[id(p)]
                part = partitions[gi]
                n = u.numels[i]
                dst = full[u.offsets[i]:u.offsets[i] + n]
                if self.stage >= 3:
                    per = _cdiv(n, self.world)
                    a, b = self.rank * per, min((self.rank + 1) * per, n)
                    if b > a:
                        dst[a:b].copy_(part[off:off + (b - a)])
                else:
                    P = self.partition_size[gi]
                    lo, hi = self.rank * P, (self.rank + 1) * P
                    a, b = max(off, lo), min(off + n, hi)
                    if a < b and part.numel() >= b - lo:
                        dst[a - off:b - off].copy_(part[a - lo:b - lo])
                    elif a < b:  # last rank's stripped padding
                        avail = part.numel() - (a - lo)
                        if avail > 0:
                            dst[a - off:a - off + avail].copy_(part[a - lo:a - lo + avail])
            if u.world > 1:
                dist.all_reduce(full, group=u.dp_group)
            s, e = u.shard_range()
            write_shard(u, full[s:e])
            del full
