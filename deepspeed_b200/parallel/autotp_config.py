"""Declarative AutoTP partitioning (reference ``module_inject/autotp_config.py``): ``TPLayerSpec`` describes, by
regular expressions over *parameter* names, how a weight is sharded (column / row / skip, fused sub-parameters through a
logical ``shape``); ``AutoTPConfig`` is an ordered list of specs (first match wins) and ``AutoTPPresets`` holds the
common model families. ``AutoTP`` (``parallel/tp.py``) consumes it when ``tensor_parallel.partition_config`` uses this
schema (``layer_specs`` / ``use_default_specs``)."""
import re
from dataclasses import dataclass, field, replace
from enum import Enum
from typing import List, Optional, Tuple, Union


class PartitionType(Enum):
    COLUMN = "column"  # shard the output dim; all-reduce in backward
    ROW = "row"  # shard the input dim; all-reduce in forward
    SKIP = "skip"  # leave replicated


def _as_tuple(x):
    return tuple(_as_tuple(v) if isinstance(v, (list, tuple)) else v for v in x) if x is not None else None


@dataclass
class TPLayerSpec:
    patterns: List[str]
    partition_type: PartitionType = PartitionType.COLUMN
    # logical view of the weight; -1 = inferred; a tuple at the partition dim lists unequal fused sub-parameters:
    #   (3, -1) three equal sub-params, ((q, k, v), -1) unequal ones, (n_experts, -1, hidden) a 3-D view
    shape: Optional[Tuple[Union[int, Tuple[int, ...]], ...]] = None
    partition_dim: Optional[int] = None  # default 0 for COLUMN, 1 for ROW
    model_types: Optional[List[str]] = None
    gather_output: bool = False  # COLUMN: all-gather the output so every rank sees the full result
    grad_allreduce: bool = False  # SKIP: replicated parameter whose gradient is summed over the TP ranks
    split_input: bool = False  # ROW: the input arrives whole (after a gathered column layer) and is sliced locally

    def __post_init__(self):
        if isinstance(self.partition_type, str):
            self.partition_type = PartitionType(self.partition_type.lower())
        self.shape = _as_tuple(self.shape)
        if self.shape is not None:
            nested = [d for d in self.shape if isinstance(d, tuple)]
            if len(nested) > 1:
                raise ValueError(f"shape {self.shape}: at most one dimension may list sub-parameter sizes")
            if any(isinstance(v, tuple) for d in nested for v in d):
                raise ValueError(f"shape {self.shape}: only one level of nesting is supported")
            if nested and self.shape.index(nested[0]) != self.get_partition_dim():
                raise ValueError(f"shape {self.shape}: sub-parameter sizes must sit at the partition dim "
                                 f"{self.get_partition_dim()}")
            if sum(1 for d in self.shape if d == -1) > 1:
                raise ValueError(f"shape {self.shape}: only one dimension may be inferred (-1)")
        self._compiled = [re.compile(p) for p in self.patterns]

    def get_partition_dim(self):
        if self.partition_dim is not None:
            return self.partition_dim
        return 1 if self.partition_type == PartitionType.ROW else 0

    def has_unequal_sub_params(self):
        return self.shape is not None and isinstance(self.shape[self.get_partition_dim()], tuple)

    def get_sub_param_sizes(self):
        return self.shape[self.get_partition_dim()] if self.has_unequal_sub_params() else None

    def get_num_sub_params(self):
        if self.shape is None:
            return None
        d = self.shape[self.get_partition_dim()]
        if isinstance(d, tuple):
            return len(d)
        return d if isinstance(d, int) and d > 0 else None

    def matches(self, param_name, model_type=None):
        if self.model_types and model_type is not None and model_type not in self.model_types:
            return False
        return any(rx.search(param_name) for rx in self._compiled)

    def sub_param_sizes_for(self, dim_size):
        """Concrete sizes of the fused sub-parameters along the partition dim of a weight (None: not fused)."""
        if self.has_unequal_sub_params():
            sizes = list(self.get_sub_param_sizes())
            if sum(sizes) != dim_size:
                raise ValueError(f"sub-parameter sizes {sizes} do not add up to {dim_size}")
            return sizes
        n = self.get_num_sub_params()
        if n and n > 1 and len(self.shape) == 2:
            if dim_size % n:
                raise ValueError(f"{dim_size} is not divisible into {n} equal sub-parameters")
            return [dim_size // n] * n
        return None


@dataclass
class AutoTPConfig:
    tp_size: int = 1
    layer_specs: List[TPLayerSpec] = field(default_factory=list)
    embedding_partition_dim: int = 1
    lm_head_patterns: List[str] = field(default_factory=lambda: ["lm_head", "embed_out"])
    use_default_specs: bool = True  # fall back to the built-in name patterns for unmatched linears
    strict_mode: bool = False  # raise when a Linear matches nothing

    def find_matching_spec(self, param_name, model_type=None):
        for spec in self.layer_specs:
            if spec.matches(param_name, model_type):
                return spec
        return None

    @classmethod
    def from_dict(cls, d):
        specs = []
        for s in d.get("layer_specs", []):
            s = dict(s)
            pats = s.pop("patterns")
            specs.append(TPLayerSpec(patterns=[pats] if isinstance(pats, str) else list(pats), **s))
        known = {k: d[k] for k in ("tp_size", "embedding_partition_dim", "lm_head_patterns", "use_default_specs",
                                   "strict_mode") if k in d}
        return cls(layer_specs=specs, **known)


def _col(*pats, **kw):
    return TPLayerSpec(patterns=list(pats), partition_type=PartitionType.COLUMN, **kw)


def _row(*pats, **kw):
    return TPLayerSpec(patterns=list(pats), partition_type=PartitionType.ROW, **kw)


def _skip(*pats, **kw):
    return TPLayerSpec(patterns=list(pats), partition_type=PartitionType.SKIP, **kw)


class AutoTPPresets:
    """Specs of common families; patterns are matched against ``<module name>.weight``."""

    @staticmethod
    def llama():
        return AutoTPConfig(layer_specs=[
            _row(r".*\.self_attn\.o_proj\.weight$", r".*\.mlp\.down_proj\.weight$"),
            _col(r".*\.self_attn\.[qkv]_proj\.weight$", r".*\.mlp\.(gate|up)_proj\.weight$"),
        ])

    @staticmethod
    def llama_gqa(num_heads, num_kv_heads, head_dim):
        q, kv = num_heads * head_dim, num_kv_heads * head_dim
        return AutoTPConfig(layer_specs=[
            _row(r".*\.self_attn\.o_proj\.weight$", r".*\.mlp\.down_proj\.weight$"),
            _col(r".*\.self_attn\.qkv_proj\.weight$", shape=((q, kv, kv), -1), partition_dim=0),
            _col(r".*\.self_attn\.[qkv]_proj\.weight$", r".*\.mlp\.(gate|up)_proj\.weight$"),
        ])

    @staticmethod
    def bloom():
        return AutoTPConfig(layer_specs=[
            _row(r".*\.self_attention\.dense\.weight$", r".*\.mlp\.dense_4h_to_h\.weight$"),
            _col(r".*\.self_attention\.query_key_value\.weight$", r".*\.mlp\.dense_h_to_4h\.weight$"),
        ])

    @staticmethod
    def chatglm():
        return AutoTPConfig(layer_specs=[
            _row(r".*\.self_attention\.dense\.weight$", r".*\.mlp\.dense_4h_to_h\.weight$"),
            _col(r".*\.self_attention\.query_key_value\.weight$", shape=(3, -1), partition_dim=0),
            _col(r".*\.mlp\.dense_h_to_4h\.weight$", shape=(2, -1), partition_dim=0),
        ])

    @staticmethod
    def mixtral():
        return AutoTPConfig(layer_specs=[
            _skip(r".*\.block_sparse_moe\.gate\.weight$"),
            _row(r".*\.self_attn\.o_proj\.weight$", r".*\.experts\.\d+\.w2\.weight$"),
            _col(r".*\.self_attn\.[qkv]_proj\.weight$", r".*\.experts\.\d+\.w[13]\.weight$"),
        ])

    @staticmethod
    def deepseek_v2():
        return AutoTPConfig(layer_specs=[
            _skip(r".*\.mlp\.gate\.weight$", r".*\.self_attn\.(q_a_proj|kv_a_proj_with_mqa)\.weight$"),
            _row(r".*\.self_attn\.o_proj\.weight$", r".*\.down_proj\.weight$"),
            _col(r".*\.self_attn\.(q_proj|q_b_proj|kv_b_proj)\.weight$", r".*\.(gate|up)_proj\.weight$"),
        ])

    @staticmethod
    def qwen2():
        return AutoTPPresets.llama()

    @staticmethod
    def phi3():
        return AutoTPConfig(layer_specs=[
            _row(r".*\.self_attn\.o_proj\.weight$", r".*\.mlp\.down_proj\.weight$"),
            _col(r".*\.self_attn\.qkv_proj\.weight$"),  # sub-parameter sizes depend on the head counts (see AutoTP)
            _col(r".*\.mlp\.gate_up_proj\.weight$", shape=(2, -1), partition_dim=0),
        ])

    @staticmethod
    def get_preset(model_type):
        table = {"llama": AutoTPPresets.llama, "mistral": AutoTPPresets.llama, "bloom": AutoTPPresets.bloom,
                 "chatglm": AutoTPPresets.chatglm, "mixtral": AutoTPPresets.mixtral,
                 "deepseek_v2": AutoTPPresets.deepseek_v2, "deepseek_v3": AutoTPPresets.deepseek_v2,
                 "qwen2": AutoTPPresets.qwen2, "qwen3": AutoTPPresets.qwen2, "phi3": AutoTPPresets.phi3}
        fn = table.get(str(model_type).lower())
        return fn() if fn is not None else None


def merge_autotp_configs(base: AutoTPConfig, override: AutoTPConfig) -> AutoTPConfig:
    """``override``'s specs come first (first match wins); scalar settings are taken from ``override``."""
    return replace(override, layer_specs=list(override.layer_specs) + list(base.layer_specs))
