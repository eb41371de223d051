"""Sequence-parallel group accessors (reference ``runtime/sequence_parallel/parallel_state_sp.py``)."""
from ...utils import groups


def initialize_sequence_parallel(sequence_parallel_size):
    groups.initialize(sp_size=sequence_parallel_size, force=True)


def get_sequence_parallel_group():
    return groups._get_sequence_parallel_group()


def get_sequence_parallel_world_size():
    return groups._get_sequence_parallel_world_size()


def get_sequence_parallel_rank():
    return groups._get_sequence_parallel_rank()


def get_sequence_data_parallel_group():
    return groups._get_sequence_data_parallel_group()


def get_sequence_data_parallel_world_size():
    return groups._get_sequence_data_parallel_world_size()


def get_sequence_data_parallel_rank():
    return groups._get_sequence_data_parallel_rank()
