"""``compile`` ds_config section (reference ``compile/config.py``); the model lives with the other config sections."""
from ..config.sections import CompileConfig  # noqa: F401
