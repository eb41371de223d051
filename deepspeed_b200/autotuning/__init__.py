from .autotuner import Autotuner  # noqa: F401
