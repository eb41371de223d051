from ..loss_scaler import *  # noqa: F401,F403
from ..loss_scaler import CreateLossScaler, DynamicLossScaler, LossScaler, LossScalerBase  # noqa: F401
