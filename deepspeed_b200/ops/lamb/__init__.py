from ..optimizers import FusedLamb
