"""Reference import path ``deepspeed.runtime.sequence_parallel`` (``ulysses_sp.py``, ``parallel_state_sp.py``); the
implementation lives in ``deepspeed_b200.sequence``."""
from . import parallel_state_sp, ulysses_sp  # noqa: F401
