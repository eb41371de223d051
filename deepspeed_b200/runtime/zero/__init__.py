"""ZeRO: flat-unit sharding (stages 0-3), construct-time partitioning, gather contexts."""
from .init_context import (Init, GatheredParameters, register_external_parameter, unregister_external_parameter,
                           is_zero_param, shutdown_init_context, restore_init_context)
from .sharded_optimizer import ZeroShardedOptimizer
from .units import FlatUnit
from ...config.sections import DeepSpeedZeroConfig, ZeroStageEnum


class MiCS_Init(Init):
    """MiCS flavour of ``Init`` (reference ``runtime/zero/mics.py:63``): shard inside sub-groups of
    ``mics_shard_size`` ranks. Partitioning at construction is identical; the shard/replica split is applied
    by the optimizer (``mics_shard_size`` in ``zero_optimization``)."""
    pass


from ...utils.z3_leaf_module import (set_z3_leaf_modules, unset_z3_leaf_modules, z3_leaf_module,  # noqa: F401,E402
                                      get_z3_leaf_modules, set_z3_leaf_module, set_z3_leaf_modules_by_name,
                                      set_z3_leaf_modules_by_suffix, z3_leaf_parameter)

from .mics import (DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3, MiCS_AllGatherCoalescedHandle,  # noqa: F401,E402
                   MiCS_CommGroups, MiCS_Offload, MiCS_Optimizer, create_mics_comm_groups)
from .handles import (AllGatherCoalescedHandle, AllGatherHandle, AllReduceCoalescedHandle, InflightParamRegistry,  # noqa: F401,E402
                      MultipleAllGatherHandles, NoGatherCoalescedHandle, NoGatherHandle, QuantizationInfo,
                      ZeROOrderedDict, ZeRoTraceMode)
from .coordinator import UnitCoordinator  # noqa: F401,E402

PartitionedParameterCoordinator = UnitCoordinator  # the reference's name for the fetch / release scheduler
from .mem_estimate import (estimate_zero2_model_states_mem_needs, estimate_zero2_model_states_mem_needs_all_cold,  # noqa: F401,E402
                           estimate_zero2_model_states_mem_needs_all_live, estimate_zero3_model_states_mem_needs,
                           estimate_zero3_model_states_mem_needs_all_cold,
                           estimate_zero3_model_states_mem_needs_all_live)
from .tiling import TiledLinear, TiledLinearReturnBias  # noqa: F401,E402
from .param_status import (DeepSpeedTensorOverride, DeepSpeedZeRoOffload, ZeroParamStatus, ZeroParamType,  # noqa: F401,E402
                           unwrap_model_for_generation)
from ...ops.quantizer import CUDAQuantizer  # noqa: F401,E402


def print_rank_0(message, debug=False, force=False):
    from ... import comm as _dist
    if (debug or force) and (not _dist.is_initialized() or _dist.get_rank() == 0):
        print(message)


def debug_rank0(msg):
    from ...utils.logging import logger
    from ... import comm as _dist
    if not _dist.is_initialized() or _dist.get_rank() == 0:
        logger.debug(msg)


def get_all_subclasses(cls, include_root=True):
    out, todo = ([cls] if include_root else []), [cls]
    while todo:
        for sub in todo.pop().__subclasses__():
            if sub not in out:
                out.append(sub)
                todo.append(sub)
    return set(out)


def free_param(param):
    """Drop the gathered copy of a ZeRO-3 parameter (its shard stays)."""
    import torch
    assert not getattr(param, "ds_active_sub_modules", None), "parameter is still in use by a running sub-module"
    param.data = torch.empty(0, dtype=param.dtype, device=param.device)
    param.ds_status = ZeroParamStatus.NOT_AVAILABLE
