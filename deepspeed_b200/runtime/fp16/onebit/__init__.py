from ...onebit import OnebitAdam, OnebitLamb, ZeroOneAdam  # noqa: F401
