from ..config.base import DeepSpeedConfigModel


class CompileConfig(DeepSpeedConfigModel):
    """``compile`` ds_config section (reference ``compile/config.py``)."""
    deepcompile: bool = False
    free_activation: bool = False
    offload_activation: bool = False
    offload_opt_states: bool = False
    double_buffer: bool = True
    symmetric_memory: bool = False
    debug_log: bool = False
    offload_parameters: bool = False
    sync_before_reduce: bool = False
    sync_after_reduce: bool = False
    sync_before_allgather: bool = False
    sync_after_allgather: bool = False
    keep_int_input_tensors: bool = True
    keep_all_input_tensors: bool = False
