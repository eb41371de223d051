from ..optimizers import FusedAdam, multi_tensor_l2norm
from ..cpu_optimizers import DeepSpeedCPUAdam
from .zenflow import ZenFlowCPUAdam, ZenFlowSelectiveAdamW, ZenFlowSelectiveAdamW_stage3  # noqa: F401,E402
