"""Sequence parallelism: Ulysses all-to-all attention, ALST tiling, FPDT chunked/offloaded attention."""
from .layer import DistributedAttention, UlyssesAttention, seq_to_heads, heads_to_seq, single_all_to_all
from .cross_entropy import vocab_sequence_parallel_cross_entropy
from .autosp_detector import detect_model_sp_info, SPModelInfo  # noqa: F401
from .autosp_vit import UlyssesSPViTAttention  # noqa: F401
from .autosp_fusion import (ModalityFusionSPAdapter, LlavaFusionAdapter, InternVLFusionAdapter,  # noqa: F401
                            Qwen2VLFusionAdapter)
from .auto_sp import auto_wrap_model_for_sp  # noqa: F401


def enable_ulysses(model, sp_group=None):
    """Make a :class:`deepspeed_b200.models.llama.LlamaForCausalLM` sequence parallel over ``sp_group``."""
    from ..utils import groups
    if sp_group is None:
        sp_group = groups._get_sequence_parallel_group()
    inner = getattr(model, "model", model)
    inner.sp_attention = UlyssesAttention(sp_group)
    return model
