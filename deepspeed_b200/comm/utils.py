"""Helpers around the communication front end (reference ``comm/utils.py``): rank / world size as the launcher
exported them, and introspection of a collective call (which argument carries the data, how many bytes it moves, how to
label it in the comms log)."""
import inspect
import os

import torch

_RANK_VARS = ("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "SLURM_PROCID")
_SIZE_VARS = ("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS")
_LOCAL_VARS = ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID")
_TENSOR_ARGS = ("tensor", "tensors", "input_list", "input_tensor_list", "input_tensor", "input")


def _first_env(names, default):
    for n in names:
        if os.environ.get(n) is not None:
            return int(os.environ[n])
    return default


def get_local_rank_from_launcher():
    return _first_env(_LOCAL_VARS, 0)


def get_world_rank_from_launcher():
    return _first_env(_RANK_VARS, 0)


def get_world_size_from_launcher():
    size = _first_env(_SIZE_VARS, 1)
    if get_world_rank_from_launcher() == 0:
        print(f"set world size to {size}")
    return size


def get_default_args(func):
    return {k: v.default for k, v in inspect.signature(func).parameters.items() if v.default is not inspect.Parameter.empty}


def get_tensor_position(func):
    """Index of the argument of a collective that holds the data (-1: none, e.g. ``barrier``)."""
    params = list(inspect.signature(func).parameters)
    for name in _TENSOR_ARGS:
        if name in params:
            return params.index(name)
    return -1


def get_tensor_kwarg(func, kwargs):
    merged = {**get_default_args(func), **kwargs}
    for name in _TENSOR_ARGS:
        if name in merged:
            return merged[name]
    return None


def _nbytes(x):
    if torch.is_tensor(x):
        return x.element_size() * x.nelement()
    if isinstance(x, (list, tuple)):
        return sum(_nbytes(t) for t in x)
    return 0


def get_msg_size_from_args(func, *args, **kwargs):
    """Bytes the call moves: the data argument may be positional, a keyword, or absent."""
    data = None
    pos = get_tensor_position(func)
    if args and 0 <= pos < len(args):
        data = args[pos]
    if data is None and kwargs:
        data = get_tensor_kwarg(func, kwargs)
    return _nbytes(data)


def get_debug_log_name(func_args, debug):
    if not debug:
        return func_args["log_name"]
    caller = "?"
    for frame in inspect.stack()[2:8]:
        if "comm" not in os.path.basename(frame.filename):
            caller = frame.function
            break
    return f"{func_args['log_name']} | [Caller Func: {caller}]"
