"""AutoSP: turn a HuggingFace model into a Ulysses sequence-parallel one without touching its code (reference
``sequence/auto_sp.py:42`` + the compile-time pass ``compile/passes/sp_compile.py`` that shards the sequence dimension
of the inputs and inserts all-to-alls around SDPA).

Here it is done eagerly: the attention function HF dispatches to is replaced by the Ulysses wrapper
(:class:`UlyssesSPAttentionHF`), the dataloader adapter shards batches on the sequence dimension, and position ids are
offset per rank - the same end state the FX passes produce, without tracing."""
from ..utils import groups
from .ulysses_sp import UlyssesSPAttentionHF, UlyssesSPDataLoaderAdapter


def auto_wrap_model_for_sp(model, sp_group=None, core_attn_implementation=None, seq_length_is_variable=True,
                           max_length=None, micro_batch_size=1):
    """Register the Ulysses attention for ``model`` (a HF ``PreTrainedModel`` or its name/path) and return the model
    parallel unit object the caller can keep (exposes ``sp_group``, ``sp_world_size``, ``sp_rank``)."""
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
