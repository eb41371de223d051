from ...config.sections import ZenFlowConfig  # noqa: F401
