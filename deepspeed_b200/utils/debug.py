"""Name look-ups and rank-aware printing for debugging (reference ``utils/debug.py``)."""
import fcntl
import os

from .. import comm as dist

module_names = {}
param_names = {}


def debug_clear_module_and_param_names():
    module_names.clear()
    param_names.clear()


def debug_extract_module_and_param_names(model):
    module_names.update({m: n for n, m in model.named_modules()})
    param_names.update({p: n for n, p in model.named_parameters()})


def debug_module2name(module):
    return module_names.get(module, "unknown")


def debug_module2name_id(module):
    return f"name={debug_module2name(module)} id={getattr(module, 'id', id(module))}"


def debug_module2name_class(module):
    return f"name={debug_module2name(module)} {module.__class__.__name__}"


def debug_param2name(param):
    return param_names.get(param, "unknown")


def ds_id(param):
    u = getattr(param, "ds_unit", None)
    return getattr(param, "ds_id", f"{u.uid}" if u is not None else "none")


def ds_shape(param):
    return tuple(getattr(param, "ds_shape", param.shape))


def debug_param2name_id(param):
    return f"name={debug_param2name(param)} id={ds_id(param)}"


def debug_param2name_id_shape(param):
    return f"name={debug_param2name(param)} id={ds_id(param)} shape={ds_shape(param)}"


def debug_param2name_id_shape_device(param):
    return f"{debug_param2name_id_shape(param)} device={param.device}"


def debug_param2name_id_numel(param):
    return f"name={debug_param2name(param)} id={ds_id(param)} numel={getattr(param, 'ds_numel', param.numel())}"


def debug_param2name_id_shape_status(param):
    return f"{debug_param2name_id_shape(param)} status={getattr(param, 'ds_status', 'n/a')}"


def printflock(*msgs):
    """print() under an exclusive file lock so the lines of different ranks do not interleave."""
    with open(__file__, "r") as fh:
        fcntl.flock(fh, fcntl.LOCK_EX)
        try:
            print(*msgs)
        finally:
            fcntl.flock(fh, fcntl.LOCK_UN)


_fh = None


def log_rank_file(rank, *msgs):
    """Append to ``log_rank_{rank}.txt`` (one file per rank: diff them to find where ranks diverge)."""
    global _fh
    if _fh is None:
        _fh = open(f"log_rank_{rank}.txt", "w")
    for m in msgs:
        _fh.write(f"{m}\n")
    _fh.flush()
    os.fsync(_fh.fileno())


def print_backward_tensors(tensor):

    def walk(fn, depth=0):
        if fn is None:
            return
        print("  " * depth + str(fn))
        for nxt, _ in getattr(fn, "next_functions", ()):
            if nxt is not None:
                t = getattr(nxt, "variable", None)
                if t is not None:
                    print("  " * (depth + 1) + f"tensor {tuple(t.shape)} requires_grad={t.requires_grad}")
                walk(nxt, depth + 1)

    walk(getattr(tensor, "grad_fn", None))


def print_rank(*msg, force=False):
    if force:
        print(f"rank {dist.get_rank() if dist.is_initialized() else 0}:", *msg)


def print_rank0(*msg, force=False):
    if force and (not dist.is_initialized() or dist.get_rank() == 0):
        print(*msg)
