from ...onebit import OnebitAdam  # noqa: F401
