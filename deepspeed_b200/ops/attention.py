"""Native flash-attention front-end (tcgen05 kernel when built; see csrc/attention_*.cu)."""
import torch

from . import kernels as K


def supported(q, k, v):
    """True when the hand-written sm_100a attention kernel can serve these operands."""
    lib = K._LIB if K._LIB is not None else None
    return bool(lib is not None and hasattr(lib, "dsb200_flash_attn_fwd") and q.dtype == torch.bfloat16
                and q.shape[-1] == 128 and _enabled())


def _enabled():
    import os
    return os.environ.get("DSB200_NATIVE_ATTN", "0") == "1"


def flash_attention(q, k, v, causal=True):
    raise NotImplementedError("native attention kernel not built in this revision")
