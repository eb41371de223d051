"""Tensor-parallel training setup (reference ``runtime/tensor_parallel/{tp_manager,config,init_utils}.py``)."""
from ...config.sections import TPConfig, TPTrainingConfig  # noqa: F401
from ...parallel.tp_plan_converter import get_hf_tp_plan as _get_hf_tp_plan  # noqa: F401
from .tp_manager import TpTrainingManager  # noqa: F401
