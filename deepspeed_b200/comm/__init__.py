"""Communication front-end (torch.distributed over NCCL/gloo) + symmetric-memory fused collectives."""
from .comm import *  # noqa: F401,F403
from .comm import (ReduceOp, ProcessGroup, comms_logger, init_distributed, is_initialized, get_rank, get_world_size,
                   get_local_rank, all_reduce, all_gather, all_gather_into_tensor, reduce_scatter_tensor,
                   all_to_all_single, broadcast, barrier, new_group, log_summary, configure, timed_op)
