"""Topology facts the writer election needs (reference ``runtime/model_checkpointing/utils.py``)."""
import os
from dataclasses import dataclass
from typing import Optional

from ... import comm as dist
from .data_parallel_writer_factory import DataParallelWriterFactory

SOCKETS_PER_MACHINE = 2


@dataclass
class MPUInfo:
    pp_world_size: int
    pp_rank: int
    tp_world_size: int
    tp_rank: int
    dp_world_size: int
    dp_peer_ranks: list
    dp_rank: int


@dataclass
class ExpertParallelInfo:
    ep_world_size: int
    ep_rank: int
    dp_world_size: int
    dp_peer_ranks: list
    dp_rank: int


@dataclass
class UniversalParallelInfo:
    global_world_size: int
    global_rank: int
    local_rank: int
    mpu_info: Optional[MPUInfo]
    ep_info: Optional[ExpertParallelInfo]
    pure_dp: bool
    num_machines: int
    machine_rank: int
    num_sockets: int


def _model_parallel_info(groups):
    """From the framework's own group state (an external mpu feeds the same state in ``groups.initialize``)."""
    tp, pp = groups._get_model_parallel_world_size(), groups._get_pipe_parallel_world_size()
    if tp * pp == 1:
        return None
    peers = dist.get_all_ranks_from_group(groups._get_data_parallel_group())
    return MPUInfo(pp, groups._get_pipe_parallel_rank(), tp, groups._get_model_parallel_rank(),
                   groups._get_data_parallel_world_size(), list(peers), groups._get_data_parallel_rank())


def _expert_parallel_info(groups):
    name = groups._get_max_expert_size_name()
    return ExpertParallelInfo(groups._get_expert_parallel_world_size(name), groups._get_expert_parallel_rank(name),
                              groups._get_expert_data_parallel_world_size(name),
                              list(groups._get_expert_data_parallel_group_ranks(name)),
                              groups._get_expert_data_parallel_rank(name))


def create_universal_parallel_info(groups, has_moe_layers):
    world = dist.get_world_size()
    local_size = int(os.environ.get("LOCAL_SIZE", os.environ.get("LOCAL_WORLD_SIZE", world)) or world)
    machines = int(os.environ.get("CROSS_SIZE", max(1, world // max(1, local_size))))
    mpu_info = _model_parallel_info(groups)
    return UniversalParallelInfo(global_world_size=world, global_rank=dist.get_rank(),
                                 local_rank=int(os.environ.get("LOCAL_RANK", 0)), mpu_info=mpu_info,
                                 ep_info=_expert_parallel_info(groups) if has_moe_layers else None,
                                 pure_dp=mpu_info is None, num_machines=machines,
                                 machine_rank=int(os.environ.get("CROSS_RANK", dist.get_rank() // max(1, local_size))),
                                 num_sockets=machines * SOCKETS_PER_MACHINE)


def create_data_parallel_writer_config(groups, parallel_unit, zero_stage, has_moe_layers):
    info = create_universal_parallel_info(groups, has_moe_layers)
    return DataParallelWriterFactory(info, parallel_unit).create_config(zero_stage, has_moe_layers)
