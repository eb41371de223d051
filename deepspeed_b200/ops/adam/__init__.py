from ..optimizers import FusedAdam, multi_tensor_l2norm
from ..cpu_optimizers import DeepSpeedCPUAdam
