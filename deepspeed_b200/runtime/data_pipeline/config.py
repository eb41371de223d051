"""Parse the ``data_efficiency`` section with defaults (reference ``runtime/data_pipeline/config.py:11``)."""
import copy


def _merge(defaults, given):
    out = copy.deepcopy(defaults)
    for k, v in (given or {}).items():
        if isinstance(v, dict) and isinstance(out.get(k), dict):
            out[k] = _merge(out[k], v)
        else:
            out[k] = v
    return out


_DEFAULTS = {
    "enabled": False,
    "seed": 1234,
    "data_sampling": {
        "enabled": False,
        "num_epochs": 1000,
        "num_workers": 0,
        "pin_memory": False,
        "curriculum_learning": {"enabled": False},
        "dynamic_batching": {
            "enabled": False, "lr_scaling_method": "linear", "min_batch_size": 1, "max_batch_size": None,
            "sequence_picking_order": "dataloader", "verbose": False
        },
    },
    "data_routing": {
        "enabled": False,
        "random_ltd": {"enabled": False, "layer_token_lr_schedule": {"enabled": False}},
    },
}


def get_data_efficiency_config(param_dict):
    out = _merge(_DEFAULTS, param_dict.get("data_efficiency", {}))
    cl = out["data_sampling"]["curriculum_learning"]
    if cl.get("enabled"):
        assert "curriculum_metrics" in cl, "Curriculum learning is enabled, 'curriculum_metrics' must be specified"
    return out


def get_data_efficiency_enabled(param_dict):
    return bool(param_dict.get("data_efficiency", {}).get("enabled", False))


def get_data_efficiency_seed(param_dict):
    return param_dict.get("data_efficiency", {}).get("seed", 1234)


def get_data_sampling(param_dict):
    return get_data_efficiency_config(param_dict)["data_sampling"]


def get_data_routing(param_dict):
    return get_data_efficiency_config(param_dict)["data_routing"]


def get_curriculum_learning(param_dict):
    return get_data_sampling(param_dict)["curriculum_learning"]


def get_curriculum_enabled_legacy(param_dict):
    return bool(param_dict.get("curriculum_learning", {}).get("enabled", False))


def get_curriculum_params_legacy(param_dict):
    if "curriculum_learning" in param_dict:
        p = copy.copy(param_dict["curriculum_learning"])
        p.pop("enabled", None)
        return p
    return False


def get_random_ltd(param_dict):
    return get_data_routing(param_dict)["random_ltd"]
