from ...onebit import ZeroOneAdam  # noqa: F401
