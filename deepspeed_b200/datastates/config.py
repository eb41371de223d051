import copy

DATASTATES_CHECKPOINTING = "datastates_ckpt"
DATASTATES_CHECKPOINTING_ENABLED = False


class DeepSpeedDataStatesConfig:
    """``datastates_ckpt`` section (reference ``datastates/config.py``): asynchronous multi-level checkpointing through
    the external DataStates-LLM library."""

    def __init__(self, param_dict):
        self.enabled = param_dict.get(DATASTATES_CHECKPOINTING, DATASTATES_CHECKPOINTING_ENABLED) is not False
        self.config = copy.deepcopy(param_dict.get(DATASTATES_CHECKPOINTING, None))
