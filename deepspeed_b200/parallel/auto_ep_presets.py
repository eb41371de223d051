"""AutoEP model presets, user configuration and validation.

Reference: ``module_inject/auto_ep_presets/{base,registry,mixtral,qwen3_moe,qwen3_5_moe,deepseek_v2,deepseek_v3}.py``
and ``module_inject/auto_ep_config.py`` (``parse_autoep_config``, ``validate_autoep_config``,
``validate_autoep_post_detection``, ``fill_autoep_config_from_hf``).

A preset says where a model family keeps its MoE block, router, routed experts and shared experts, how the experts are
stored (one fused 3-D parameter per projection - Transformers >= 5 - or a ``ModuleList`` of MLPs) and how the router
scores; an adapter adds the family-specific rules (DeepSeek group-limited routing read from the HF config, layout
fallback to ``ModuleList`` experts, which forward contract the replacement block must honour).
"""
import copy
from dataclasses import dataclass, field, replace
from typing import Any, Dict, List, Optional, Tuple

import torch.nn as nn

from ..utils.logging import logger

_UNSET = object()  # "key absent": differs from None, which for ``expert_w3`` means "gate and up are fused in w1"


def _reject_load_balance_coeff(value):
    raise ValueError(f"load_balance_coeff={value!r} is not supported (auxiliary-loss-free balancing through a dynamic "
                     "expert bias); set it to null or omit the key")


@dataclass
class MoEModelPreset:
    moe_layer_pattern: str
    router_pattern: str
    experts_pattern: str
    expert_storage: str  # "fused_3d" | "module_list"
    expert_w1: str
    expert_w2: str
    expert_w3: Optional[str]
    num_experts_attr: str
    top_k_attr: str
    score_func: str  # "softmax" | "sigmoid"
    score_apply: str  # "pre" | "post"
    route_norm: bool
    gate_bias: bool
    has_shared_experts: bool = False
    shared_experts_pattern: str = ""
    shared_experts_gate_pattern: str = ""
    autoep_config_defaults: Dict[str, Any] = field(default_factory=dict)
    supports_expert_bias: bool = True
    unsupported_router_bias_names: Tuple[str, ...] = ()
    preset_adapter: str = "default"
    hf_model_types: Tuple[str, ...] = ()
    unsupported_hf_model_type_notes: Dict[str, str] = field(default_factory=dict)
    min_transformers_version: Optional[str] = None
    validated_transformers_versions: str = ""
    docs_support_notes: str = ""


@dataclass
class MoELayerSpec:
    moe_module_name: str
    model_family: str
    router_name: str
    experts_name: str
    expert_storage: str
    expert_w1_name: str
    expert_w2_name: str
    expert_w3_name: Optional[str]
    num_experts: int
    top_k: int
    hidden_size: int
    ffn_hidden_size: int
    score_func: str
    score_apply: str
    route_norm: bool
    gate_bias: bool
    return_router_logits: bool
    router_logits_capture_target: str = "none"
    router_logits_capture_index: Optional[int] = None
    router_logits_capture_layer_name: Optional[str] = None
    has_shared_experts: bool = False
    shared_experts_name: str = ""
    shared_experts_gate_name: str = ""
    route_scale: float = 1.0
    num_expert_groups: Optional[int] = None
    num_limited_groups: Optional[int] = None
    group_score_func: str = "top2_sum"
    supports_expert_bias: bool = True
    unsupported_router_bias_names: Tuple[str, ...] = ()
    preset_adapter: str = "default"
    router_logits_capture_mode: str = "post_score"
    moe_output_shape: str = "batched"


@dataclass
class AutoEPConfig:
    enabled: bool = False
    autoep_size: int = 1
    expert_tensor_parallel_size: int = 1
    validate_folding_routing: bool = False
    preset_model: Optional[str] = None
    moe_layer_pattern: Optional[str] = None
    expert_pattern: Optional[str] = None
    router_pattern: Optional[str] = None
    use_grouped_mm: bool = True
    disable_triton_grouped_mm: bool = False
    route_norm: Optional[bool] = None
    route_scale: float = 1.0
    score_apply: str = "auto"
    combine_impl: str = "auto"
    num_expert_groups: Optional[int] = None
    num_limited_groups: Optional[int] = None
    score_func: str = "auto"
    top_k: Any = "auto"
    load_balance_coeff: Any = _UNSET
    routed_scaling_factor: Any = "auto"
    expert_w1: Optional[str] = None
    expert_w2: Optional[str] = None
    expert_w3: Any = _UNSET
    num_experts_attr: Optional[str] = None
    top_k_attr: Optional[str] = None
    has_shared_experts: Optional[bool] = None
    shared_experts_pattern: Optional[str] = None
    shared_experts_gate_pattern: Optional[str] = None
    return_router_logits: Optional[bool] = None  # None: decided by the preset adapter
    _load_balance_coeff_explicit: bool = field(default=False, init=False, repr=False)

    def __post_init__(self):
        self._load_balance_coeff_explicit = self.load_balance_coeff is not _UNSET
        if self.load_balance_coeff is _UNSET:
            self.load_balance_coeff = None


@dataclass(frozen=True)
class GroupRoutingConfig:
    num_expert_groups: Optional[int]
    num_limited_groups: Optional[int]
    group_score_func: str = "top2_sum"


@dataclass(frozen=True)
class ForwardContract:
    return_router_logits: bool = False
    capture_target: str = "none"  # "moe_block" | "router" | "none"
    capture_index: Optional[int] = None
    capture_layer_name: Optional[str] = None
    router_logits_capture_mode: str = "post_score"  # "raw" | "post_score"
    moe_output_shape: str = "batched"  # "batched" | "flat"


def _version_tuple(v):
    out = []
    for part in str(v).split("."):
        digits = "".join(ch for ch in part if ch.isdigit())
        if digits == "":
            break
        out.append(int(digits))
    if not out:
        raise ValueError(f"unparsable version {v!r}")
    return tuple(out)


class AutoEPPresetAdapter:
    """Family rules shared by every preset; subclasses override what differs."""
    checks_transformers_version = False

    def installed_transformers_version(self):
        import transformers
        return getattr(transformers, "__version__", "unknown")

    def validate_compatibility(self, preset_name, preset, model_config):
        model_type = getattr(model_config, "model_type", None) if model_config is not None else None
        if model_type is None:
            return
        note = preset.unsupported_hf_model_type_notes.get(model_type)
        if note is not None:
            ok = ", ".join(repr(t) for t in preset.hf_model_types) or "none"
            raise ValueError(f"AutoEP preset '{preset_name}' does not support model_type='{model_type}'. {note} "
                             f"Supported HF model_type value(s): {ok}.")
        need = preset.min_transformers_version
        if need is None or not self.checks_transformers_version or model_type not in preset.hf_model_types:
            return
        try:
            have = self.installed_transformers_version()
        except Exception as e:  # noqa: BLE001
            raise ValueError(f"AutoEP preset '{preset_name}' for model_type='{model_type}' requires Transformers >= "
                             f"{need}, but transformers could not be imported: {e}.") from e
        try:
            too_old = _version_tuple(have) < _version_tuple(need)
        except ValueError as e:
            raise ValueError(f"AutoEP preset '{preset_name}' for model_type='{model_type}' requires Transformers >= "
                             f"{need}, but the installed version '{have}' could not be parsed.") from e
        if too_old:
            raise ValueError(f"AutoEP preset '{preset_name}' for model_type='{model_type}' requires Transformers >= "
                             f"{need}, but installed transformers=={have}.")

    def resolve_route_norm(self, config, preset, model_config):
        if config.route_norm is not None:
            return config.route_norm
        norm = getattr(model_config, "norm_topk_prob", None)
        return bool(norm) if norm is not None else preset.route_norm

    def resolve_group_routing(self, config, model_config):
        return GroupRoutingConfig(config.num_expert_groups, config.num_limited_groups)

    def resolve_expert_layout(self, experts_module, preset):
        """Fall back to the ``ModuleList`` layout when the fused default does not match what the model holds."""
        if not isinstance(experts_module, nn.ModuleList) or len(experts_module) == 0:
            return preset
        if not (preset.expert_storage == "fused_3d" and preset.expert_w3 is None):
            return preset
        e0 = experts_module[0]
        for names in (("gate_proj", "down_proj", "up_proj"), ("w1", "w2", "w3")):
            if all(hasattr(getattr(e0, n, None), "weight") for n in names):
                return replace(preset, expert_storage="module_list", expert_w1=names[0], expert_w2=names[1],
                               expert_w3=names[2])
        return preset

    def adjust_forward_contract(self, contract, moe_module=None, expert_storage="fused_3d"):
        return contract


class RouterHookAdapter(AutoEPPresetAdapter):
    """Mixtral / Qwen-MoE: Transformers >= 5 records router logits with a forward hook on the router module, and the
    MoE block returns one tensor. The replacement keeps the original router module in place (``router.gate``) and fires
    its hooks, so nothing has to be re-registered; the ``ModuleList`` layout of older releases returned a
    ``(hidden, router_logits)`` pair from the block instead."""
    checks_transformers_version = True

    def adjust_forward_contract(self, contract, moe_module=None, expert_storage="fused_3d"):
        if expert_storage == "module_list":
            return replace(contract, return_router_logits=True, capture_target="moe_block", capture_index=1,
                           router_logits_capture_mode="raw")
        return replace(contract, return_router_logits=False, capture_target="router", capture_index=0,
                       router_logits_capture_mode="raw")


class DeepSeekV2PresetAdapter(AutoEPPresetAdapter):
    """Keeps the native (un-normalised) top-k weights; group-limited greedy routing scores a group by its best expert."""
    checks_transformers_version = True

    def resolve_route_norm(self, config, preset, model_config):
        return config.route_norm if config.route_norm is not None else preset.route_norm

    def resolve_group_routing(self, config, model_config):
        base = super().resolve_group_routing(config, model_config)
        if getattr(model_config, "topk_method", None) != "group_limited_greedy":
            return base
        return GroupRoutingConfig(base.num_expert_groups or getattr(model_config, "n_group", None),
                                  base.num_limited_groups or getattr(model_config, "topk_group", None), "max")


class DeepSeekV3PresetAdapter(AutoEPPresetAdapter):
    """Group-limited routing always on when the HF config carries ``n_group`` / ``topk_group`` (top-2 sum per group)."""
    checks_transformers_version = True

    def resolve_group_routing(self, config, model_config):
        base = super().resolve_group_routing(config, model_config)
        return GroupRoutingConfig(base.num_expert_groups or getattr(model_config, "n_group", None),
                                  base.num_limited_groups or getattr(model_config, "topk_group", None),
                                  base.group_score_func)


class Qwen35MoePresetAdapter(RouterHookAdapter):
    """Qwen3.5 text backbone: the router records (logits, scores, indices); hooks read index 1 of the block output on
    older layouts."""

    def adjust_forward_contract(self, contract, moe_module=None, expert_storage="fused_3d"):
        c = super().adjust_forward_contract(contract, moe_module, expert_storage)
        return replace(c, capture_target="router")


_COMMON = dict(moe_layer_pattern=r"model\.layers\.\d+\.mlp", router_pattern="gate", experts_pattern="experts",
               expert_storage="fused_3d", expert_w1="gate_up_proj", expert_w2="down_proj", expert_w3=None,
               top_k_attr="num_experts_per_tok", score_apply="post", gate_bias=False, min_transformers_version="5.0.0")

PRESET_MODELS: Dict[str, MoEModelPreset] = {
    "mixtral":
    MoEModelPreset(**_COMMON, num_experts_attr="num_local_experts", score_func="softmax", route_norm=True,
                   preset_adapter="router_hook", hf_model_types=("mixtral", )),
    "qwen3_moe":
    MoEModelPreset(**_COMMON, num_experts_attr="num_experts", score_func="softmax", route_norm=True,
                   has_shared_experts=True, shared_experts_pattern="shared_expert",
                   shared_experts_gate_pattern="shared_expert_gate", preset_adapter="router_hook",
                   hf_model_types=("qwen3_moe", "qwen2_moe"),
                   docs_support_notes="Also covers Qwen2-MoE (shared expert behind a sigmoid gate)."),
    "qwen3_5_moe":
    MoEModelPreset(**{**_COMMON, "min_transformers_version": "5.2.0"}, num_experts_attr="num_experts",
                   score_func="softmax", route_norm=True, has_shared_experts=True,
                   shared_experts_pattern="shared_expert", shared_experts_gate_pattern="shared_expert_gate",
                   preset_adapter="qwen3_5_moe", hf_model_types=("qwen3_5_moe_text", ),
                   unsupported_hf_model_type_notes={
                       "qwen3_5_moe": "AutoEP converts the Qwen3.5 text backbone; pass the text model/config "
                                      "(model_type='qwen3_5_moe_text')."
                   }),
    "deepseek_v2":
    MoEModelPreset(**_COMMON, num_experts_attr="n_routed_experts", score_func="softmax", route_norm=False,
                   has_shared_experts=True, shared_experts_pattern="shared_experts",
                   autoep_config_defaults={"load_balance_coeff": None}, supports_expert_bias=False,
                   preset_adapter="deepseek_v2", hf_model_types=("deepseek_v2", )),
    "deepseek_v3":
    MoEModelPreset(**_COMMON, num_experts_attr="n_routed_experts", score_func="sigmoid", route_norm=False,
                   has_shared_experts=True, shared_experts_pattern="shared_experts",
                   autoep_config_defaults={"load_balance_coeff": None}, supports_expert_bias=False,
                   preset_adapter="deepseek_v3", hf_model_types=("deepseek_v3", )),
}

_PRESET_ADAPTERS: Dict[str, AutoEPPresetAdapter] = {
    "default": AutoEPPresetAdapter(),
    "router_hook": RouterHookAdapter(),
    "qwen3_5_moe": Qwen35MoePresetAdapter(),
    "deepseek_v2": DeepSeekV2PresetAdapter(),
    "deepseek_v3": DeepSeekV3PresetAdapter(),
}


def register_preset(name, preset, adapter=None):
    """Add a model family at run time (``adapter``: instance registered under ``preset.preset_adapter``)."""
    if adapter is not None:
        _PRESET_ADAPTERS[preset.preset_adapter] = adapter
    if preset.preset_adapter not in _PRESET_ADAPTERS:
        raise RuntimeError(f"preset '{name}' names the unregistered adapter '{preset.preset_adapter}'")
    PRESET_MODELS[name] = preset


def available_preset_names():
    return tuple(PRESET_MODELS)


def get_preset(name):
    if name not in PRESET_MODELS:
        raise ValueError(f"Unknown preset_model '{name}'. Available presets: {list(PRESET_MODELS)}")
    return PRESET_MODELS[name]


def get_preset_adapter(name):
    if name not in _PRESET_ADAPTERS:
        raise ValueError(f"Unknown AutoEP preset adapter '{name}'")
    return _PRESET_ADAPTERS[name]


def preset_name_for_hf_model_type(model_type):
    return next((n for n, p in PRESET_MODELS.items() if model_type in p.hf_model_types), None)


def resolve_autoep_config_defaults(config, preset_name):
    """Preset-level defaults for the keys the user left out (today: ``load_balance_coeff``)."""
    defaults = PRESET_MODELS[preset_name].autoep_config_defaults if preset_name in PRESET_MODELS else {}
    if not defaults:
        return config
    out = copy.copy(config)
    if "load_balance_coeff" in defaults and not config._load_balance_coeff_explicit:
        out.load_balance_coeff = defaults["load_balance_coeff"]
    return out


_OVERRIDES = (("moe_layer_pattern", "moe_layer_pattern"), ("router_pattern", "router_pattern"),
              ("expert_pattern", "experts_pattern"), ("expert_w1", "expert_w1"), ("expert_w2", "expert_w2"),
              ("num_experts_attr", "num_experts_attr"), ("top_k_attr", "top_k_attr"),
              ("has_shared_experts", "has_shared_experts"), ("shared_experts_pattern", "shared_experts_pattern"),
              ("shared_experts_gate_pattern", "shared_experts_gate_pattern"))


def apply_config_overrides(config, preset):
    """A copy of ``preset`` with the user's explicit fields applied (the registered preset is never modified)."""
    over = {dst: getattr(config, src) for src, dst in _OVERRIDES if getattr(config, src) is not None}
    if config.expert_w3 is not _UNSET:
        over["expert_w3"] = config.expert_w3
    return replace(preset, **over) if over else preset


def _custom_preset(config):
    return MoEModelPreset(moe_layer_pattern=config.moe_layer_pattern, router_pattern=config.router_pattern or "gate",
                          experts_pattern=config.expert_pattern or "experts", expert_storage="fused_3d",
                          expert_w1=config.expert_w1 or "gate_up_proj", expert_w2=config.expert_w2 or "down_proj",
                          expert_w3=None if config.expert_w3 is _UNSET else config.expert_w3,
                          num_experts_attr=config.num_experts_attr or "num_local_experts",
                          top_k_attr=config.top_k_attr or "num_experts_per_tok",
                          score_func=config.score_func if config.score_func != "auto" else "softmax",
                          score_apply=config.score_apply if config.score_apply != "auto" else "post",
                          route_norm=config.route_norm if config.route_norm is not None else True, gate_bias=False,
                          has_shared_experts=bool(config.has_shared_experts),
                          shared_experts_pattern=config.shared_experts_pattern or "",
                          shared_experts_gate_pattern=config.shared_experts_gate_pattern or "")


def resolve_preset_candidates(config, model_config) -> List[Tuple[str, MoEModelPreset]]:
    """Ordered presets to try: the named one, else the one matching the HF ``model_type``, else a custom pattern,
    else every registered preset."""

    def checked(name):
        p = apply_config_overrides(config, get_preset(name))
        get_preset_adapter(p.preset_adapter).validate_compatibility(name, p, model_config)
        return [(name, p)]

    if config.preset_model is not None:
        return checked(config.preset_model)
    model_type = getattr(model_config, "model_type", None) if model_config is not None else None
    if model_type:
        name = preset_name_for_hf_model_type(model_type)
        if name is not None:
            return checked(name)
        for name, p in PRESET_MODELS.items():
            if model_type in p.unsupported_hf_model_type_notes:
                get_preset_adapter(p.preset_adapter).validate_compatibility(name, p, model_config)
    if config.moe_layer_pattern:
        return [("custom", _custom_preset(config))]
    return [(n, apply_config_overrides(config, p)) for n, p in PRESET_MODELS.items()]


# ---- ds_config section ------------------------------------------------------------------------------------------
def parse_autoep_config(param_dict) -> AutoEPConfig:
    """``expert_parallel`` section of the ds_config (a dict or an already parsed section object)."""
    if not param_dict:
        return AutoEPConfig()
    if not isinstance(param_dict, dict):
        d = param_dict.model_dump() if hasattr(param_dict, "model_dump") else dict(vars(param_dict))
        fields_set = getattr(param_dict, "model_fields_set", None)
        if fields_set is not None:
            d = {k: v for k, v in d.items() if k in fields_set or k in ("enabled", "autoep_size")}
        param_dict = d
    known = {f for f in AutoEPConfig.__dataclass_fields__ if not f.startswith("_")}
    kw = {k: v for k, v in param_dict.items() if k in known and k not in ("load_balance_coeff", "expert_w3")}
    cfg = AutoEPConfig(**kw)
    if "load_balance_coeff" in param_dict:
        if param_dict["load_balance_coeff"] is not None:
            _reject_load_balance_coeff(param_dict["load_balance_coeff"])
        cfg._load_balance_coeff_explicit = True
    if "expert_w3" in param_dict:
        cfg.expert_w3 = param_dict["expert_w3"]
    return cfg


def _check_attr_name(name, value):
    if value is None:
        return
    if not isinstance(value, str) or value == "":
        raise ValueError(f"{name} must be a non-empty string")
    if "." in value:
        raise ValueError(f"{name} must be a direct attribute name (no dots)")


def validate_autoep_config(config, world_size, pp_size=1, tp_size=1, sp_size=1, *, zero_stage=0,
                           deepcompile_enabled=False, tp_preset_model=None, use_data_before_expert_parallel=False,
                           mpu=None, zero_offload_optimizer=False, zero_offload_param=False):
    """Raises ``ValueError`` on a configuration that cannot run (sizes that do not tile the world, unknown preset or
    enum value, group-routing bounds, attribute names with dots, ...)."""
    from .auto_ep_folding import build_folding_spec, validate_folding_global
    if config.load_balance_coeff is not None:
        _reject_load_balance_coeff(config.load_balance_coeff)
    if not isinstance(config.validate_folding_routing, bool):
        raise ValueError("expert_parallel.validate_folding_routing must be a boolean")
    if not config.enabled:
        return
    spec = build_folding_spec(world_size=world_size, pp_size=pp_size, tp_size=max(tp_size, 1),
                              ep_size=config.autoep_size, etp_size=config.expert_tensor_parallel_size,
                              mp_mode="tp" if tp_size > 1 else "sp")
    validate_folding_global(spec, zero_stage=zero_stage, sp_size=sp_size, deepcompile_enabled=deepcompile_enabled,
                            use_data_before_expert_parallel=use_data_before_expert_parallel, mpu=mpu,
                            tp_preset=tp_preset_model, ep_preset=config.preset_model,
                            zero_offload_optimizer=zero_offload_optimizer, zero_offload_param=zero_offload_param)
    if config.preset_model is not None and config.preset_model not in PRESET_MODELS:
        raise ValueError(f"Unknown preset_model '{config.preset_model}'. Available presets: {list(PRESET_MODELS)}")
    for name, allowed in (("score_apply", ("auto", "pre", "post")),
                          ("combine_impl", ("auto", "weighted_sum", "legacy_bmm")),
                          ("score_func", ("auto", "softmax", "sigmoid"))):
        if getattr(config, name) not in allowed:
            raise ValueError(f"{name} must be one of {allowed}, got '{getattr(config, name)}'")
    if config.num_limited_groups is not None and config.num_limited_groups < 1:
        raise ValueError(f"num_limited_groups must be >= 1, got {config.num_limited_groups}")
    if config.num_expert_groups is not None:
        if config.num_expert_groups < 1:
            raise ValueError(f"num_expert_groups must be >= 1, got {config.num_expert_groups}")
        if config.num_limited_groups is not None and config.num_limited_groups > config.num_expert_groups:
            raise ValueError(f"num_limited_groups ({config.num_limited_groups}) must be <= num_expert_groups "
                             f"({config.num_expert_groups})")
    if config.autoep_size == 1:
        logger.warning("autoep_size=1: every rank owns all experts, MoE blocks are replaced but nothing is exchanged")
    for name in ("expert_w1", "expert_w2", "num_experts_attr", "top_k_attr", "router_pattern", "expert_pattern",
                 "shared_experts_pattern", "shared_experts_gate_pattern"):
        _check_attr_name(name, getattr(config, name))
    if config.expert_w3 is not _UNSET and config.expert_w3 is not None:
        _check_attr_name("expert_w3", config.expert_w3)
    if config.has_shared_experts is not None and not isinstance(config.has_shared_experts, bool):
        raise ValueError("has_shared_experts must be a boolean when set")
    if config.routed_scaling_factor != "auto" and (isinstance(config.routed_scaling_factor, bool)
                                                   or not isinstance(config.routed_scaling_factor, (int, float))):
        raise ValueError("routed_scaling_factor must be a number or 'auto'")
    if config.has_shared_experts is True and not config.shared_experts_pattern:
        logger.warning("has_shared_experts=True without shared_experts_pattern: shared experts cannot be located")
    if config.shared_experts_pattern and config.has_shared_experts is False:
        logger.warning(f"shared_experts_pattern='{config.shared_experts_pattern}' is ignored: has_shared_experts=False")


def _divisors(n):
    return [d for d in range(1, n + 1) if n % d == 0]


def validate_autoep_post_detection(config, specs):
    """Checks that need the detected layers: the expert count against ``autoep_size`` and the routing groups."""
    if not config.enabled or not specs:
        return
    for s in specs:
        if config.autoep_size > s.num_experts:
            raise ValueError(f"autoep_size={config.autoep_size} exceeds num_experts={s.num_experts} in layer "
                             f"'{s.moe_module_name}': each rank must own at least one expert. Valid autoep_size "
                             f"values: {_divisors(s.num_experts)}")
        if s.num_experts % config.autoep_size != 0:
            raise ValueError(f"num_experts={s.num_experts} in layer '{s.moe_module_name}' is not divisible by "
                             f"autoep_size={config.autoep_size}. Suggested autoep_size values: "
                             f"{_divisors(s.num_experts)}")
        g = s.num_expert_groups if s.num_expert_groups is not None else config.num_expert_groups
        lim = s.num_limited_groups if s.num_limited_groups is not None else config.num_limited_groups
        if lim is not None and g is None:
            raise ValueError(f"num_limited_groups requires num_expert_groups in layer '{s.moe_module_name}'")
        if g is not None:
            if g < 1 or s.num_experts % g != 0:
                raise ValueError(f"num_expert_groups ({g}) must be >= 1 and divide num_experts ({s.num_experts}) in "
                                 f"layer '{s.moe_module_name}'")
            if lim is None:
                raise ValueError(f"num_limited_groups must be set when num_expert_groups is set in layer "
                                 f"'{s.moe_module_name}'")
            if not 1 <= lim <= g:
                raise ValueError(f"num_limited_groups ({lim}) must be in [1, num_expert_groups={g}] in layer "
                                 f"'{s.moe_module_name}'")


def fill_autoep_config_from_hf(config, model_config):
    """Take DeepSeek-style routing fields the user did not repeat in the ds_config from the HF model config."""
    if model_config is None:
        return
    # DeepSeek-V2 checkpoints carry n_group / topk_group even when ``topk_method`` is plain greedy top-k
    grouped = getattr(model_config, "topk_method", "group_limited_greedy") not in ("greedy", None)
    if grouped and (getattr(model_config, "n_group", None) or 0) > 1:
        if config.num_expert_groups is None:
            config.num_expert_groups = getattr(model_config, "n_group", None)
        if config.num_limited_groups is None:
            config.num_limited_groups = getattr(model_config, "topk_group", None)
    if config.routed_scaling_factor == "auto":
        s = getattr(model_config, "routed_scaling_factor", None)
        if s is not None:
            config.route_scale = float(s)
    else:
        config.route_scale = float(config.routed_scaling_factor)
