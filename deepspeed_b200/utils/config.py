"""``timers`` config section helpers under the reference module path (``utils/config.py``)."""
from ..config.base import DeepSpeedConfigModel

TIMERS = "timers"
TIMERS_THROUGHPUT = "throughput"


class DeepSpeedThroughputTimerConfig(DeepSpeedConfigModel):
    enabled: bool = True
    synchronized: bool = True  # synchronize the device around the measurement (accurate, slower)


def get_timers_config(param_dict):
    section = (param_dict or {}).get(TIMERS, {}).get(TIMERS_THROUGHPUT, {})
    return DeepSpeedThroughputTimerConfig(**section)
