"""Shape classification of fused gate/up expert tensors (reference: ``moe/fused_expert_layout.py``).

Transformers >= 5 stores the routed experts of a block as two 3-D parameters. Either ``gate_up_proj [E, 2I, H]`` with
``down_proj [E, H, I]`` (``F.linear`` orientation, what :class:`GroupedExperts` uses), or the transposed
``[E, H, 2I]`` / ``[E, I, H]`` (``x @ w`` orientation)."""
from dataclasses import dataclass
from typing import Optional, Tuple


@dataclass(frozen=True)
class FusedExpertLayout:
    layout: str  # "gate_up_first" | "hidden_first"
    hidden_size: int
    ffn_hidden_size: int
    needs_transpose: bool


def classify_fused_gate_up_layout(w1_shape: Tuple[int, ...], w2_shape: Tuple[int, ...]) -> Optional[FusedExpertLayout]:
    if len(w1_shape) != 3 or len(w2_shape) != 3 or w1_shape[0] != w2_shape[0]:
        return None
    _, a, b = w1_shape
    if a % 2 == 0 and tuple(w2_shape[1:]) == (b, a // 2):
        return FusedExpertLayout("gate_up_first", b, a // 2, False)
    if b % 2 == 0 and tuple(w2_shape[1:]) == (b // 2, a):
        return FusedExpertLayout("hidden_first", a, b // 2, True)
    return None
