"""Normalise the ``compression_training`` section: fill defaults for every technique's ``shared_parameters`` and
``different_groups`` (reference ``compression/config.py:11``)."""
import copy

from .constants import *  # noqa: F401,F403

_SHARED_DEFAULTS = {
    WEIGHT_QUANTIZATION: {
        "enabled": False, "quantizer_kernel": False, "schedule_offset": 0, "quantize_groups": 1,
        "quantize_verbose": False, "quantization_type": "symmetric", "quantize_weight_in_forward": False,
        "rounding": "nearest", "fp16_mixed_quantize": {"enabled": False, "quantize_change_ratio": 0.001},
    },
    ACTIVATION_QUANTIZATION: {
        "enabled": False, "quantization_type": "symmetric", "range_calibration": "dynamic", "schedule_offset": 1000,
    },
    SPARSE_PRUNING: {
        "enabled": False, "method": "l1", "schedule_offset": 1000, "block_pattern": "4x1", "dense_ratio": 0.1,
        "excluded_modules": [], "schedule_offset_stride": 1,
    },
    ROW_PRUNING: {"enabled": False, "method": "l1", "schedule_offset": 1000},
    HEAD_PRUNING: {"enabled": False, "method": "topk", "schedule_offset": 1000},
    CHANNEL_PRUNING: {"enabled": False, "method": "l1", "schedule_offset": 1000},
}

_REQUIRED_GROUP_PARAMS = {
    WEIGHT_QUANTIZATION: ("start_bits", "target_bits"),
    ACTIVATION_QUANTIZATION: ("bits", ),
    SPARSE_PRUNING: (),
    ROW_PRUNING: ("dense_ratio", ),
    HEAD_PRUNING: ("dense_ratio", ),
    CHANNEL_PRUNING: ("dense_ratio", ),
}


def _technique(name, given, require_groups=True):
    given = given or {}
    shared = copy.deepcopy(_SHARED_DEFAULTS[name])
    user_shared = given.get(SHARED_PARAMETERS, {})
    for k, v in user_shared.items():
        if isinstance(v, dict) and isinstance(shared.get(k), dict):
            shared[k].update(v)
        else:
            shared[k] = v
    if name == WEIGHT_QUANTIZATION:
        mixed = shared.pop("fp16_mixed_quantize")
        shared["fp16_mixed_quantize"] = mixed["enabled"]
        shared["quantize_change_ratio"] = mixed["quantize_change_ratio"]
    if name == SPARSE_PRUNING:
        shared.setdefault("schedule_offset_end", shared["schedule_offset"])
    if name == HEAD_PRUNING and shared["enabled"]:
        assert "num_heads" in shared, "head_pruning requires shared_parameters.num_heads"
    groups = {}
    for gname, g in given.get(DIFFERENT_GROUPS, {}).items():
        assert DIFFERENT_GROUPS_PARAMETERS in g, f"{name}/{gname}: '{DIFFERENT_GROUPS_PARAMETERS}' is required"
        params = dict(g[DIFFERENT_GROUPS_PARAMETERS])
        for req in _REQUIRED_GROUP_PARAMS[name]:
            assert req in params, f"{name}/{gname}: params.{req} is required"
        if name == WEIGHT_QUANTIZATION:
            params.setdefault("quantization_period", 1)
        if name == SPARSE_PRUNING and shared["method"] != "snip_momentum":
            assert "dense_ratio" in params, f"{name}/{gname}: params.dense_ratio is required"
        groups[gname] = {
            DIFFERENT_GROUPS_PARAMETERS: params,
            DIFFERENT_GROUPS_MODULE_SCOPE: g.get(DIFFERENT_GROUPS_MODULE_SCOPE, DIFFERENT_GROUPS_MODULE_SCOPE_DEFAULT),
            DIFFERENT_GROUPS_RELATED_MODULE_SCOPE: g.get(DIFFERENT_GROUPS_RELATED_MODULE_SCOPE,
                                                         DIFFERENT_GROUPS_RELATED_MODULE_SCOPE_DEFAULT),
        }
    if shared["enabled"] and require_groups:
        assert groups, f"{name} is enabled, {DIFFERENT_GROUPS} must be specified"
    return {SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: groups}


def get_layer_reduction(section):
    lr = dict(section.get(LAYER_REDUCTION, {}))
    lr.setdefault(LAYER_REDUCTION_ENABLED, False)
    return lr


def get_compression_config(param_dict):
    section = param_dict.get(COMPRESSION_TRAINING, {}) or {}
    out = {LAYER_REDUCTION: get_layer_reduction(section)}
    for name in TECHNIQUES:
        out[name] = _technique(name, section.get(name))
    return out


# ---- per-technique readers under the reference's names (``compression/config.py:65-452``) -------------------------
# Each takes the ``compression_training`` section (``get_<technique>``) or one technique's sub-dict
# (``..._shared_parameters`` / ``..._different_groups``) and returns the normalised dictionaries built above.
def _whole(name, param_dict):
    return _technique(name, (param_dict or {}).get(name))


def _shared(name, sub_param_dict):
    given = {SHARED_PARAMETERS: (sub_param_dict or {}).get(SHARED_PARAMETERS, {})}
    return _technique(name, given, require_groups=False)[SHARED_PARAMETERS]


def _groups(name, sub_param_dict):
    sub = sub_param_dict or {}
    shared = dict(sub.get(SHARED_PARAMETERS, {}))
    shared["enabled"] = False  # only the groups are wanted here: no "groups must exist" check
    if name == HEAD_PRUNING:
        shared.setdefault("num_heads", 0)
    return _technique(name, {SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: sub.get(DIFFERENT_GROUPS, {})})[DIFFERENT_GROUPS]


def get_weight_quantization(param_dict):
    return _whole(WEIGHT_QUANTIZATION, param_dict)


def get_weight_quantization_shared_parameters(param_dict):
    return _shared(WEIGHT_QUANTIZATION, param_dict)


def get_weight_quantization_different_groups(param_dict):
    return _groups(WEIGHT_QUANTIZATION, param_dict)


def get_activation_quantization(param_dict):
    return _whole(ACTIVATION_QUANTIZATION, param_dict)


def get_activation_quantization_shared_parameters(param_dict):
    return _shared(ACTIVATION_QUANTIZATION, param_dict)


def get_activation_quantization_different_groups(param_dict):
    return _groups(ACTIVATION_QUANTIZATION, param_dict)


def get_sparse_pruning(param_dict):
    return _whole(SPARSE_PRUNING, param_dict)


def get_sparse_pruning_shared_parameters(param_dict):
    return _shared(SPARSE_PRUNING, param_dict)


def get_sparse_pruning_different_groups(param_dict):
    return _groups(SPARSE_PRUNING, param_dict)


def get_row_pruning(param_dict):
    return _whole(ROW_PRUNING, param_dict)


def get_row_pruning_shared_parameters(param_dict):
    return _shared(ROW_PRUNING, param_dict)


def get_row_pruning_different_groups(param_dict):
    return _groups(ROW_PRUNING, param_dict)


def get_head_pruning(param_dict):
    return _whole(HEAD_PRUNING, param_dict)


def get_head_pruning_shared_parameters(param_dict):
    return _shared(HEAD_PRUNING, param_dict)


def get_head_pruning_different_groups(param_dict):
    return _groups(HEAD_PRUNING, param_dict)


def get_channel_pruning(param_dict):
    return _whole(CHANNEL_PRUNING, param_dict)


def get_channel_pruning_shared_parameters(param_dict):
    return _shared(CHANNEL_PRUNING, param_dict)


def get_channel_pruning_different_groups(param_dict):
    return _groups(CHANNEL_PRUNING, param_dict)


def get_layer_reduction_enabled(param_dict):
    return bool(get_layer_reduction(param_dict or {}).get(LAYER_REDUCTION_ENABLED, False))


def get_layer_reduction_params(param_dict):
    lr = get_layer_reduction(param_dict or {})
    lr.pop(LAYER_REDUCTION_ENABLED, None)
    return lr if LAYER_REDUCTION in (param_dict or {}) else False


def get_quantize_enabled(param_dict):
    """Weight quantization of the ``compression_training`` section switched on (whole ds_config given)."""
    section = (param_dict or {}).get(COMPRESSION_TRAINING)
    if not section:
        return False
    return bool(get_weight_quantization_shared_parameters(section.get(WEIGHT_QUANTIZATION, {})).get("enabled", False))
