from ...onebit import OnebitLamb  # noqa: F401
