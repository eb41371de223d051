from ..optimizers import FusedAdagrad
from ..cpu_optimizers import DeepSpeedCPUAdagrad
