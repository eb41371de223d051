"""DeepSpeed-MoE: capacity-based (GShard) MoE layer + dropless AutoEP experts."""
from .layer import MoE
from .sharded_moe import TopKGate, MOELayer, top1gating, top2gating, topkgating
from .experts import Experts
