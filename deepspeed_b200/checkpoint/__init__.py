"""Checkpointing: reference-layout save/load, offline fp32 consolidation, universal checkpoints."""
from .zero_to_fp32 import (convert_zero_checkpoint_to_fp32_state_dict, get_fp32_state_dict_from_zero_checkpoint,
                           load_state_dict_from_zero_checkpoint)
