"""AutoSP: turn a HuggingFace model into a Ulysses sequence-parallel one without touching its code (reference
``sequence/auto_sp.py:42`` + the compile-time pass ``compile/passes/sp_compile.py`` that shards the sequence dimension
of the inputs and inserts all-to-alls around SDPA).

Here it is done eagerly: the attention function HF dispatches to is replaced by the Ulysses wrapper
(:class:`UlyssesSPAttentionHF`), the dataloader adapter shards batches on the sequence dimension, and position ids are
offset per rank - the same end state the FX passes produce, without tracing."""
from ..utils import groups
from .ulysses_sp import UlyssesSPAttentionHF, UlyssesSPDataLoaderAdapter


def wrap_vit_attention_for_sp(model, process_group):
    """Reference ``auto_wrap_model_for_sp(model, process_group)``: replace every recognised ViT attention module by
    :class:`UlyssesSPViTAttention` in place; LLM attention is served by the HF attention hook below, the projection
    layer by a :class:`ModalityFusionSPAdapter` the caller attaches. Raises when nothing is recognised."""
    from ..utils.logging import logger
    from .autosp_detector import VIT_ATTENTION, detect_model_sp_info
    from .autosp_vit import UlyssesSPViTAttention
    info = detect_model_sp_info(model)
    if not info.vit_attn_modules and not info.llm_attn_modules:
        raise ValueError("auto_wrap_model_for_sp: no recognisable attention modules found. Add the model's attention "
                         "class name(s) to VIT_ATTENTION / LLM_ATTENTION in deepspeed/sequence/autosp_detector.py")
    for name, module in info.vit_attn_modules:
        parent = model
        parts = name.split(".")
        for part in parts[:-1]:
            parent = getattr(parent, part)
        setattr(parent, parts[-1], UlyssesSPViTAttention(module, process_group,
                                                         has_cls_token=VIT_ATTENTION.get(type(module).__name__, True)))
    logger.info(f"AutoSP: wrapped {len(info.vit_attn_modules)} ViT attention layer(s)")
    if info.llm_attn_modules:
        logger.info(f"AutoSP: {len(info.llm_attn_modules)} LLM attention layer(s) use the HuggingFace attention "
                    "interface; enable Ulysses for them with UlyssesSPAttentionHF.register_with_transformers")
    if info.vision_projection_module is not None:
        logger.warning(f"AutoSP: wrap the vision projection '{info.vision_projection_module[0]}' with the "
                       "ModalityFusionSPAdapter subclass of your architecture")
    return model


def auto_wrap_model_for_sp(model, sp_group=None, core_attn_implementation=None, seq_length_is_variable=True,
                           max_length=None, micro_batch_size=1, process_group=None):
    """Two call forms. ``auto_wrap_model_for_sp(module, process_group)`` (reference signature): wrap the ViT attention
    modules of a multimodal ``nn.Module`` in place and return it. Otherwise: register the Ulysses attention for
    ``model`` (a HF ``PreTrainedModel`` or its name/path) and return the model parallel unit object the caller can
    keep (exposes ``sp_group``, ``sp_world_size``, ``sp_rank``)."""
    import torch
    pg = process_group if process_group is not None else sp_group
    if isinstance(model, torch.nn.Module) and pg is not None and core_attn_implementation is None:
        from .autosp_detector import detect_model_sp_info
        if detect_model_sp_info(model).vit_attn_modules or process_group is not None:
            return wrap_vit_attention_for_sp(model, pg)
    sp_group = sp_group if sp_group is not None else groups._get_sequence_parallel_group()
    return UlyssesSPAttentionHF.register_with_transformers(
        model_name_or_path=model, core_attn_implementation=core_attn_implementation or "sdpa",
        sequence_parallel_size=groups._get_sequence_parallel_world_size(), max_length=max_length,
        micro_batch_size=micro_batch_size, seq_length_is_variable=seq_length_is_variable)


def shard_dataloader_for_sp(dataloader, device=None):
    """Wrap a dataloader so that every SP rank receives its contiguous slice of each sequence."""
    return UlyssesSPDataLoaderAdapter(dataloader, sp_rank=groups._get_sequence_parallel_rank(),
                                      sp_group=groups._get_sequence_parallel_group(),
                                      sp_world_size=groups._get_sequence_parallel_world_size(), device=device)


class AutoSPConfig:

    def __init__(self, enabled=False, sequence_parallel_size=1, attn_implementation="sdpa"):
        self.enabled, self.sequence_parallel_size = enabled, sequence_parallel_size
        self.attn_implementation = attn_implementation
