"""MiCS: ZeRO-3 sharding inside sub-groups of the data-parallel world, replication across them.

Reference: ``runtime/zero/mics.py`` (``MiCS_Init``, ``MiCS_Optimizer``, ``MiCS_Offload``,
``MiCS_AllGatherCoalescedHandle``) and ``runtime/zero/mics_utils.py`` (``MiCS_CommGroups``,
``create_mics_comm_groups``). Parameters are partitioned over ``mics_shard_size`` ranks only, so a gather crosses at
most that many ranks (one NVSwitch domain when the shard size is the node size); gradients are reduce-scattered inside
the shard group and all-reduced across the replica groups. The optimizer side of this lives in
``ZeroShardedOptimizer`` (``mics_shard_size`` argument); this module holds the topology math, the hierarchical
all-gather for shard groups that span nodes, and the reference's class names."""
import os
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from ... import comm as dist
from .sharded_optimizer import ZeroShardedOptimizer


def mics_rank_tables(world_size, shard_size, ndev_per_node=None):
    """Rank lists without process groups: consecutive ``shard_size`` ranks shard one copy of the model; ranks with the
    same position in their shard group hold the same shard (replica group). When a shard group spans several nodes
    the hierarchical gather uses, per shard group, one inter-node group per local device index and one intra-node
    group per node."""
    assert shard_size >= 1 and world_size % shard_size == 0, \
        f"mics_shard_size={shard_size} must divide the data-parallel world size {world_size}"
    shard = [list(range(i, i + shard_size)) for i in range(0, world_size, shard_size)]
    repl = [list(range(p, world_size, shard_size)) for p in range(shard_size)]
    out = {"shard_groups": shard, "replicate_groups": repl, "span_nodes": 1, "inter_node_groups": [],
           "intra_node_groups": []}
    if ndev_per_node and shard_size > ndev_per_node:
        assert shard_size % ndev_per_node == 0, "a shard group must cover whole nodes"
        out["span_nodes"] = shard_size // ndev_per_node
        for g in shard:
            out["intra_node_groups"] += [g[i:i + ndev_per_node] for i in range(0, shard_size, ndev_per_node)]
            out["inter_node_groups"] += [g[d::ndev_per_node] for d in range(ndev_per_node)]
    return out


@dataclass
class MiCS_CommGroups:
    param_shard_group: object = None
    param_shard_size: int = -1
    param_shard_rank: int = -1
    param_repli_group: object = None
    param_repli_size: int = -1
    param_repli_rank: int = -1
    param_intra_node_group: object = None
    param_inter_node_shard_group: object = None
    shard_ranks: List[int] = field(default_factory=list)
    ndev_per_node: Optional[int] = None


def create_mics_comm_groups(shard_size, dp_group=None, hierarchical_allgather=False, mpu=None, ndev_per_node=None):
    """Collective: every rank creates every group. ``NDEV_PER_NODE`` overrides the devices-per-node count."""
    ranks = dist.get_all_ranks_from_group(dp_group) if dp_group is not None else list(range(dist.get_world_size()))
    if mpu is not None:
        assert dp_group is None or dp_group == mpu.get_data_parallel_group()
    ndev = int(os.environ.get("NDEV_PER_NODE", ndev_per_node or max(1, torch.cuda.device_count())))
    t = mics_rank_tables(len(ranks), shard_size, ndev if hierarchical_allgather else None)
    me = dist.get_rank()
    g = MiCS_CommGroups(ndev_per_node=ndev)

    def make(tables):
        mine, mine_ranks = None, None
        for rl in tables:
            glob = [ranks[i] for i in rl]
            pg = dist.new_group(glob) if len(glob) > 1 else None
            if me in glob:
                mine, mine_ranks = pg, glob
        return mine, mine_ranks

    g.param_shard_group, g.shard_ranks = make(t["shard_groups"])
    g.param_shard_size, g.param_shard_rank = len(g.shard_ranks), g.shard_ranks.index(me)
    rg, rr = make(t["replicate_groups"])
    g.param_repli_group, g.param_repli_size, g.param_repli_rank = rg, len(rr), rr.index(me)
    if hierarchical_allgather and t["span_nodes"] > 1:
        g.param_intra_node_group, _ = make(t["intra_node_groups"])
        g.param_inter_node_shard_group, _ = make(t["inter_node_groups"])
    return g


def hierarchical_all_gather(output, shard, groups: MiCS_CommGroups):
    """All-gather of a shard group that spans nodes in two steps: among the ranks with the same local index on the
    other nodes first (one NIC hop per rank, all NICs in parallel), then inside the node over NVLink. ``output``
    (``shard_size * shard.numel()``) ends up in shard-rank order."""
    if groups.param_inter_node_shard_group is None or groups.param_intra_node_group is None:
        dist.all_gather_into_tensor(output, shard, group=groups.param_shard_group)
        return output
    ndev = groups.ndev_per_node
    nodes = groups.param_shard_size // ndev
    n = shard.numel()
    inter = shard.new_empty(nodes * n)  # my local index on every node: [node, n]
    dist.all_gather_into_tensor(inter, shard.contiguous().view(-1), group=groups.param_inter_node_shard_group)
    intra = shard.new_empty(ndev * nodes * n)  # [dev, node, n]
    dist.all_gather_into_tensor(intra, inter, group=groups.param_intra_node_group)
    output.view(nodes, ndev, n).copy_(intra.view(ndev, nodes, n).transpose(0, 1))
    return output


class MiCS_AllGatherCoalescedHandle:
    """Work handle of a (hierarchical) gather of several parameters: ``wait()`` completes the communication and marks
    the parameters available."""

    def __init__(self, allgather_handle, params, partitions=None, world_size=1):
        self.allgather_handle, self.params, self.partitions, self.world_size = allgather_handle, params, partitions, \
            world_size
        self.complete = False

    def wait(self, **kwargs):
        if self.complete:
            return
        if self.allgather_handle is not None and hasattr(self.allgather_handle, "wait"):
            self.allgather_handle.wait()
        from .param_status import ZeroParamStatus
        for p in self.params:
            if hasattr(p, "ds_status"):
                p.ds_status = ZeroParamStatus.AVAILABLE
        self.complete = True


class DeepSpeedZeroOptimizer(ZeroShardedOptimizer):
    """ZeRO stage 1 / 2 under the reference's class name (``runtime/zero/stage_1_and_2.py``): optimizer state (and
    with ``partition_gradients`` the gradients) sharded over the data-parallel group."""

    def __init__(self, module, optimizer, *, partition_gradients=False, **kw):
        kw.setdefault("stage", 2 if partition_gradients else 1)
        assert kw["stage"] in (1, 2)
        super().__init__(module, optimizer, **kw)


class DeepSpeedZeroOptimizer_Stage3(ZeroShardedOptimizer):
    """ZeRO stage 3 under the reference's class name (``runtime/zero/stage3.py``)."""

    def __init__(self, module, optimizer, **kw):
        kw["stage"] = 3
        super().__init__(module, optimizer, **kw)


class MiCS_Optimizer(DeepSpeedZeroOptimizer_Stage3):
    """Stage 3 with sub-group sharding; ``mics_shard_size`` is mandatory."""

    def __init__(self, module, optimizer, *, mics_shard_size, **kw):
        assert mics_shard_size and mics_shard_size > 0, "MiCS needs zero_optimization.mics_shard_size"
        super().__init__(module, optimizer, mics_shard_size=mics_shard_size, **kw)


class MiCS_Offload:
    """Parameter-offload front of MiCS for inference-only runs: stage 3 without an optimizer."""

    def __new__(cls, module, *, mics_shard_size, **kw):
        return MiCS_Optimizer(module, None, mics_shard_size=mics_shard_size, **kw)
