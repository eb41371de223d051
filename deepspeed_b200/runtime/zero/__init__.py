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


def set_z3_leaf_modules(model, leaf_module_classes):
    """Mark module classes whose children must not become separate units (reference utils/z3_leaf_module.py)."""
    found = []
    for m in model.modules():
        if isinstance(m, tuple(leaf_module_classes)):
            m._z3_leaf = True
            found.append(m)
    return found


def unset_z3_leaf_modules(model, leaf_module_classes):
    out = []
    for m in model.modules():
        if isinstance(m, tuple(leaf_module_classes)) and getattr(m, "_z3_leaf", False):
            m._z3_leaf = False
            out.append(m)
    return out


def z3_leaf_module(model):
    return getattr(model, "_z3_leaf", False)


def get_z3_leaf_modules(model):
    return [m for m in model.modules() if getattr(m, "_z3_leaf", False)]
