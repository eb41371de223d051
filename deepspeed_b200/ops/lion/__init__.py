from ..optimizers import FusedLion
from ..cpu_optimizers import DeepSpeedCPULion
