"""``comms_logger`` configuration holder (reference ``comm/config.py``)."""
from ..config.sections import CommsLoggerConfig  # noqa: F401


class DeepSpeedCommsConfig:

    def __init__(self, ds_config):
        self.comms_logger_enabled = "comms_logger" in ds_config
        if self.comms_logger_enabled:
            self.comms_logger = CommsLoggerConfig(**ds_config["comms_logger"])
