from .zenflow_config import ZenFlowConfig  # noqa: F401
from .zenflow_utils import ZenFlowController  # noqa: F401
