"""Checkpointing: reference-layout save/load, offline fp32 consolidation, universal checkpoints."""
from .zero_to_fp32 import (convert_zero_checkpoint_to_fp32_state_dict, get_fp32_state_dict_from_zero_checkpoint,
                           load_state_dict_from_zero_checkpoint)
from .reshape import (DeepSpeedCheckpoint, ZeROCheckpoint, get_layer_ckpt_name_for_rank,  # noqa: F401,E402
                      get_model_3d_descriptor, get_model_ckpt_name_for_rank, get_zero_ckpt_name_for_rank, merge_state,
                      model_3d_desc, reshape_meg_2d_parallel)
from .universal import SubparamShape, enable_universal_checkpoint  # noqa: F401,E402
from .constants import *  # noqa: F401,F403,E402
