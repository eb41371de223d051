"""Backwards-compatible queries on Megatron / DeepSpeed ``mpu`` objects (reference ``utils/bwc.py``): the tensor- and
pipeline-parallel accessors were renamed several times; try the names newest first."""


def _first(mpu, names, default):
    if mpu is None:
        return default
    for n in names:
        fn = getattr(mpu, n, None)
        if fn is not None:
            return fn()
    return default


def bwc_tensor_model_parallel_rank(mpu=None):
    return _first(mpu, ("get_tensor_model_parallel_rank", "get_slice_parallel_rank", "get_model_parallel_rank"), 0)


def bwc_tensor_model_parallel_world_size(mpu=None):
    return _first(mpu, ("get_tensor_model_parallel_world_size", "get_slice_parallel_world_size",
                        "get_model_parallel_world_size"), 1)


def bwc_tensor_model_parallel_group(mpu=None):
    return _first(mpu, ("get_tensor_model_parallel_group", "get_slice_parallel_group", "get_model_parallel_group"),
                  None)


def bwc_pipeline_parallel_world_size(mpu=None):
    return _first(mpu, ("get_pipeline_model_parallel_world_size", "get_pipe_parallel_world_size"), 1)


def bwc_pipeline_parallel_group(mpu=None):
    return _first(mpu, ("get_pipeline_model_parallel_group", "get_pipe_parallel_group"), None)
