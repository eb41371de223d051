"""Parameter status enums, the module-hook owner and the generation helper of the reference ZeRO-3 API
(``runtime/zero/partition_parameters.py:47,223,236``, ``parameter_offload.py:130``, ``stage3.py:92``)."""
import contextlib
from enum import Enum


class ZeroParamType(Enum):
    NORMAL = 1  # an ordinary torch parameter
    PARTITIONED = 2  # sharded over the data-parallel group
    REMOTE = 3  # held by a single rank


class ZeroParamStatus(str, Enum):
    """``param.ds_status``. A ``str`` enum: this framework stores the plain names, so both
    ``p.ds_status == ZeroParamStatus.AVAILABLE`` and ``p.ds_status == "AVAILABLE"`` hold."""
    AVAILABLE = "AVAILABLE"  # full parameter present on every rank
    NOT_AVAILABLE = "NOT_AVAILABLE"  # only the shard is held
    INFLIGHT = "INFLIGHT"  # gather launched, not yet waited for


class DeepSpeedTensorOverride(Enum):
    """Which tensor-constructor arguments ``zero.Init`` overrides."""
    dtype = 1
    device = 2


DEFAULT_TENSOR_OVERRIDES = [DeepSpeedTensorOverride.dtype, DeepSpeedTensorOverride.device]


class DeepSpeedZeRoOffload:
    """Owner of the ZeRO-3 module hooks. The reference registers fetch / release hooks on every sub-module
    (``parameter_offload.py:331``); here hooks exist per flat unit and belong to the ``UnitCoordinator``. This class
    gives that coordinator the reference's handle: ``forward_hooks`` / ``backward_hooks`` lists, re-registration,
    ``get_param_coordinator`` and ``empty_partition_cache``."""

    def __init__(self, optimizer=None, module=None, **unused):
        self.optimizer = optimizer
        self.module = module if module is not None else getattr(optimizer, "module", None)
        self.backward_hooks = []  # pre-backward hooks are tensor hooks created per forward: nothing persistent

    @property
    def coordinator(self):
        return getattr(self.optimizer, "coordinator", None)

    @property
    def forward_hooks(self):
        co = self.coordinator
        return co.handles if co is not None else []

    @forward_hooks.setter
    def forward_hooks(self, value):
        co = self.coordinator
        if co is not None:
            co.handles = list(value)

    def get_param_coordinator(self, training=True):
        return self.coordinator

    def empty_partition_cache(self):
        co = self.coordinator
        if co is not None:
            co.release_all()

    def _register_deepspeed_module(self, module=None):
        co = self.coordinator
        if co is not None and not co.handles:
            co._install()

    def destroy(self):
        co = self.coordinator
        if co is not None:
            co.remove_hooks()


@contextlib.contextmanager
def unwrap_model_for_generation(model):
    """Gather every ZeRO-3 parameter once and take the per-unit hooks off for the duration (token-by-token generation
    would otherwise gather and release the whole model at every step). ``model``: the engine."""
    from .init_context import GatheredParameters
    opt = getattr(model, "optimizer", None)
    co = getattr(opt, "coordinator", None)
    with GatheredParameters(list(model.parameters())):
        if co is not None:
            co.remove_hooks()
        try:
            yield model
        finally:
            if co is not None:
                co._install()
