"""ds_config: one pydantic tree for the whole configuration surface."""
from .base import DeepSpeedConfigModel, get_config_default, pp_int
from .config import (DeepSpeedConfig, DeepSpeedConfigError, DeepSpeedConfigWriter, load_ds_config)
from .sections import *  # noqa: F401,F403
