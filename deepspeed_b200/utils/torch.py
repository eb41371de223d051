"""torch version helpers (reference ``utils/torch.py``)."""
import torch
from packaging import version as _v


def required_torch_version(min_version=None, max_version=None):
    assert min_version or max_version, "Must provide a min_version or max_version argument"
    cur = _v.parse(torch.__version__)
    if min_version and _v.parse(str(min_version)) > cur:
        return False
    if max_version and _v.parse(str(max_version)) < cur:
        return False
    return True


def is_functorch_transforming():
    """True under torch.func.grad / vmap / jacrev (gradient hooks must not fire collectives there)."""
    ft = getattr(torch._C, "_functorch", None)
    peek = getattr(ft, "peek_interpreter_stack", None) if ft is not None else None
    return peek is not None and peek() is not None


def register_grad_hook(param, hook):
    """Hook that runs after the gradient was accumulated into ``param.grad``."""
    return param.register_post_accumulate_grad_hook(hook)


def jit_script_compat(fn):
    """The reference compiles small elementwise helpers with torch.compile / TorchScript; here those helpers are CUDA
    kernels, so the function is returned unchanged."""
    return fn
