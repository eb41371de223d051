"""Who writes which share of a replicated checkpoint state.

Reference: ``runtime/model_checkpointing/data_parallel_writer_factory.py``. Ranks of one data-parallel group (or
expert-data-parallel group) hold identical state, so any subset of them can share the write; the ``data_parallel``
unit says which: every ``REPLICA``, one rank per ``SOCKET``, one per ``MACHINE``. Without a unit only group rank 0
writes. ZeRO-3 shards the weights themselves: every rank writes its own file, alone.

Election rule used here: ranks of the group are bucketed by the resource they sit on (``global_rank //
ranks_per_resource``); the first rank of each bucket writes, and the writers are numbered by bucket order. Groups of
different model-parallel slices consist of different ranks, so their writers already spread over the machines without
the reference's explicit resource partitioning between slices."""
from dataclasses import dataclass

from .constants import CheckpointDataParallel


@dataclass
class DataParallelWriterConfig:
    world_size: int  # number of writers sharing this state
    rank: int  # this writer's share
    global_rank: int
    local_rank: int
    pure_dp: bool


class DataParallelWriterFactory:

    def __init__(self, uni_parallel_info, parallel_unit):
        self._info = uni_parallel_info
        self._unit = parallel_unit
        n_res = uni_parallel_info.num_sockets if parallel_unit == CheckpointDataParallel.SOCKET \
            else uni_parallel_info.num_machines
        self._num_resources = max(1, n_res)
        self._ranks_per_resource = max(1, uni_parallel_info.global_world_size // self._num_resources)

    def _config(self, world_size, rank):
        i = self._info
        return DataParallelWriterConfig(world_size, rank, i.global_rank, i.local_rank, i.pure_dp)

    def create_config(self, zero_stage, has_moe_layers):
        """This rank's writer configuration, or ``None`` when it does not write replicated state."""
        if int(zero_stage) == 3:
            return self._config(1, 0)
        i = self._info
        if has_moe_layers and i.ep_info is not None:
            peers, my = list(i.ep_info.dp_peer_ranks), i.ep_info.dp_rank
        elif i.mpu_info is not None:
            peers, my = list(i.mpu_info.dp_peer_ranks), i.mpu_info.dp_rank
        else:
            peers, my = list(range(i.global_world_size)), i.global_rank
        cfg = self._elect(peers, my)
        if cfg is None and int(zero_stage) >= 1:
            # optimizer state is partitioned over the group: every rank still writes its own shard, unshared
            return self._config(1, 0)
        return cfg

    def _elect(self, peers, my_index):
        if self._unit is None:
            return self._config(1, 0) if my_index == 0 else None
        if self._unit == CheckpointDataParallel.REPLICA:
            return self._config(len(peers), my_index)
        buckets = {}
        for r in peers:
            buckets.setdefault(r // self._ranks_per_resource, []).append(r)
        writers = [ranks[0] for _, ranks in sorted(buckets.items())]
        me = self._info.global_rank
        return self._config(len(writers), writers.index(me)) if me in writers else None
