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

from .tiling import TiledLinear, TiledLinearReturnBias  # noqa: F401,E402
from .param_status import (DeepSpeedTensorOverride, DeepSpeedZeRoOffload, ZeroParamStatus, ZeroParamType,  # noqa: F401,E402
                           unwrap_model_for_generation)
