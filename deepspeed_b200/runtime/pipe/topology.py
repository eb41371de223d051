"""Rank grids for pipeline x data (x model) parallelism and the process groups derived from them.

API follows the reference ``runtime/pipe/topology.py`` (``ProcessTopology`` :12: ``get_rank``, ``get_coord``,
``get_axis_comm_lists``, ``filter_match``, ``get_axis_list``, ``get_rank_repr``; ``PipelineParallelGrid`` :251
exposing the ``mpu`` interface). The grid is a row-major ``numpy`` index space: rank = ravel(coord).
"""
from collections import namedtuple

import numpy as np

from ... import comm as dist


class ProcessTopology:
    """Row-major mapping between named coordinates and global ranks (last axis varies fastest)."""

    def __init__(self, axes, dims):
        self.axes, self.dims = list(axes), [int(d) for d in dims]
        self.ProcessCoord = namedtuple("ProcessCoord", self.axes)
        self._grid = np.arange(int(np.prod(self.dims))).reshape(self.dims)
        self.mapping = {self.get_coord(r): r for r in range(self._grid.size)}

    def world_size(self):
        return int(self._grid.size)

    def get_axis_names(self):
        return self.axes

    def get_dim(self, axis):
        return self.dims[self.axes.index(axis)] if axis in self.axes else 0

    def get_coord(self, rank):
        if not 0 <= rank < self._grid.size:
            raise ValueError(f"rank {rank} not found in topology.")
        return self.ProcessCoord(*(int(i) for i in np.unravel_index(rank, self.dims)))

    def get_rank(self, **coord):
        if set(coord) != set(self.axes):
            raise ValueError("get_rank() does not support slices. Use filter_match())")
        idx = tuple(coord[a] for a in self.axes)
        for i, d in zip(idx, self.dims):
            assert 0 <= i < d, f"key {coord} invalid"
        return int(self._grid[idx])

    def _take(self, fixed):
        """Ranks whose coordinates match ``fixed`` (axis -> index), ascending."""
        sl = tuple(fixed.get(a, slice(None)) for a in self.axes)
        return [int(r) for r in np.atleast_1d(self._grid[sl]).reshape(-1)]

    def filter_match(self, **fixed):
        return self._take(fixed)

    def get_axis_list(self, axis, idx):
        return self._take({axis: idx})

    def get_axis_comm_lists(self, axis):
        """All rank lists that vary only along ``axis`` (one communicator per combination of the others)."""
        if axis not in self.axes:
            return []
        k = self.axes.index(axis)
        moved = np.moveaxis(self._grid, k, -1).reshape(-1, self.dims[k])
        return [[int(r) for r in row] for row in moved]

    def get_rank_repr(self, rank, omit_axes=("data", "pipe"), inner_sep="_", outer_sep="-"):
        c = self.get_coord(rank)
        return outer_sep.join(f"{a}{inner_sep}{getattr(c, a):02d}" for a in self.axes if a not in set(omit_axes))

    def __str__(self):
        return str(self.mapping)


class PipeDataParallelTopology(ProcessTopology):
    """pipe x data with data fastest: DP reductions stay between neighbouring ranks."""

    def __init__(self, num_pp, num_dp):
        super().__init__(["pipe", "data"], [num_pp, num_dp])


class PipeModelDataParallelTopology(ProcessTopology):

    def __init__(self, num_pp, num_mp, num_dp):
        super().__init__(["pipe", "data", "model"], [num_pp, num_dp, num_mp])


def _balanced_pp_dp(world):
    """Split world into pp x dp by dealing prime factors alternately (pp gets the first)."""
    pp = dp = 1
    n, f, turn = world, 2, 0
    while n > 1:
        if n % f == 0:
            if turn % 2 == 0:
                pp *= f
            else:
                dp *= f
            turn += 1
            n //= f
        else:
            f += 1
    return pp, dp


class PipelineParallelGrid:
    """Owns the process groups of a topology and answers the Megatron-style ``mpu`` queries."""

    def __init__(self, topology=None, process_group=None):
        self.global_rank, self.world_size = dist.get_rank(), dist.get_world_size()
        if topology is None:
            pp, dp = _balanced_pp_dp(self.world_size)
            topology = PipeDataParallelTopology(num_pp=pp, num_dp=dp)
        self._topo = topology
        assert topology.world_size() == self.world_size, "Invalid Grid"
        dim = lambda a: max(topology.get_dim(a), 1)  # noqa: E731
        self.data_parallel_size, self.pipe_parallel_size, self.model_parallel_size = dim("data"), dim("pipe"), dim("model")
        self.slice_parallel_size = self.model_parallel_size
        me = topology.get_coord(self.global_rank)
        self.stage_id, self.data_parallel_id = me.pipe, me.data
        self.is_first_stage, self.is_last_stage = self.stage_id == 0, self.stage_id == self.pipe_parallel_size - 1

        # "model" group in the DeepSpeed sense = every rank sharing my data-parallel index (pipe x slice)
        model_lists = [sorted(topology.get_axis_list("data", d)) for d in range(self.data_parallel_size)]
        self.ds_model_proc_group, ranks = self._groups(model_lists)
        self.ds_model_world_size, self.ds_model_rank = len(ranks), ranks.index(self.global_rank)
        self.dp_groups = topology.get_axis_comm_lists("data")
        self.dp_proc_group, self.dp_group = self._groups(self.dp_groups)
        self.pp_groups = topology.get_axis_comm_lists("pipe")
        self.pp_proc_group, self.pp_group = self._groups(self.pp_groups)
        self.p2p_groups = [[r, self._next_in_pipe(r)] for r in range(self.world_size)]
        slice_lists = topology.get_axis_comm_lists("model") or [[r] for r in range(self.world_size)]
        self.slice_proc_group, self.slice_group = self._groups(slice_lists)

    def _groups(self, rank_lists):
        mine = (None, None)
        for ranks in rank_lists:
            g = dist.new_group(ranks=list(ranks)) if self.world_size > 1 else None
            if self.global_rank in ranks:
                mine = (g, list(ranks))
        return mine

    def _next_in_pipe(self, rank):
        c = self._topo.get_coord(rank)._asdict()
        c["pipe"] = (c["pipe"] + 1) % self.pipe_parallel_size
        return self._topo.get_rank(**c)

    def stage_to_global(self, stage_id, **kwargs):
        c = self._topo.get_coord(self.global_rank)._asdict()
        c.update(pipe=stage_id, **kwargs)
        return self._topo.get_rank(**c)

    def topology(self):
        return self._topo

    def get_stage_id(self):
        return self.stage_id

    def get_data_parallel_id(self):
        return self.data_parallel_id

    # ---- mpu interface ----
    def get_global_rank(self):
        return self.global_rank

    def get_pipe_parallel_rank(self):
        return self.stage_id

    def get_pipe_parallel_world_size(self):
        return self.pipe_parallel_size

    def get_pipe_parallel_group(self):
        return self.pp_proc_group

    def get_data_parallel_rank(self):
        return self.data_parallel_id

    def get_data_parallel_world_size(self):
        return self.data_parallel_size

    def get_data_parallel_group(self):
        return self.dp_proc_group

    def get_model_parallel_rank(self):
        return self.ds_model_rank

    def get_model_parallel_world_size(self):
        return self.ds_model_world_size

    def get_model_parallel_group(self):
        return self.ds_model_proc_group

    def get_pipeline_model_parallel_rank(self):
        return self.get_pipe_parallel_rank()

    def get_pipeline_model_parallel_world_size(self):
        return self.get_pipe_parallel_world_size()

    def get_data_parallel_group_ranks(self):
        """Global ranks of this rank's data-parallel group."""
        return self.dp_group

    def get_tensor_model_parallel_rank(self):
        return self.get_slice_parallel_rank()

    def get_tensor_model_parallel_world_size(self):
        return self.get_slice_parallel_world_size()

    def get_tensor_model_parallel_group(self):
        return self.get_slice_parallel_group()

    def get_slice_parallel_rank(self):
        return getattr(self._topo.get_coord(self.global_rank), "model", 0)

    def get_slice_parallel_world_size(self):
        return self.slice_parallel_size

    def get_slice_parallel_group(self):
        return self.slice_proc_group
