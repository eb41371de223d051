"""Package-level constants (reference ``deepspeed/constants.py``)."""
import os
from datetime import timedelta

TORCH_DISTRIBUTED_DEFAULT_PORT = 29500
# process-group timeout; DEEPSPEED_TIMEOUT is in minutes
default_pg_timeout = timedelta(minutes=int(os.getenv("DEEPSPEED_TIMEOUT", default=10)))
INFERENCE_GENERIC_MODE = "generic"
INFERENCE_SPECIALIZED_MODE = "specialized"
CROSS_RANK = "CROSS_RANK"
CROSS_SIZE = "CROSS_SIZE"
LOCAL_RANK = "LOCAL_RANK"
