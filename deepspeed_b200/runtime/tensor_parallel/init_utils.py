"""Bookkeeping between ``tp_model_init`` and ``initialize`` (reference ``runtime/tensor_parallel/init_utils.py``): what
``tp_model_init`` was called with is remembered, checked for consistency across calls, and merged into the ds_config
(``deepspeed_b200.merge_tp_model_init_into_config``)."""
import base64
import json
import os


def load_ds_config(config):
    """A dict, a path to a json / hjson file, or a base64-encoded json string."""
    if isinstance(config, dict):
        return config
    if isinstance(config, str):
        if os.path.exists(config):
            with open(config) as f:
                text = f.read()
            try:
                return json.loads(text)
            except ValueError:
                import hjson
                return hjson.loads(text)
        try:
            return json.loads(base64.urlsafe_b64decode(config).decode("utf-8"))
        except Exception as exc:  # noqa: BLE001
            raise ValueError("Expected a string path to an existing deepspeed config, or a dictionary or a valid "
                             f"base64. Received: {config}") from exc
    raise ValueError(f"Expected a string path to an existing deepspeed config, or a dictionary or a valid base64. "
                     f"Received: {config}")


def tp_group_world_size(tp_group, dist_module):
    if tp_group is None or dist_module is None:
        return None
    return dist_module.get_world_size(group=tp_group)


def record_tp_model_init_args(tp_size, dtype, tp_group, dist_module):
    """Remember the arguments of ``tp_model_init``; a second call must agree with the first."""
    import deepspeed_b200 as ds
    new = {"tp_size": tp_size, "dtype": dtype, "tp_group": tp_group}
    old = ds._TP_MODEL_INIT_ARGS
    if old is None:
        ds._TP_MODEL_INIT_ARGS = new
        return
    conflict = ValueError("Conflicting tp_model_init arguments detected across multiple calls.")
    if old["tp_size"] != tp_size or old.get("dtype") != dtype:
        raise conflict
    a, b = old.get("tp_group"), tp_group
    if (a is None) != (b is None):
        raise conflict
    if a is not None and tp_group_world_size(a, dist_module) != tp_group_world_size(b, dist_module):
        raise conflict


def infer_config_dtype(config_dict):
    """``torch.bfloat16`` / ``torch.float16`` when the config enables that precision, else ``None``."""
    import torch
    for key, dt in (("bf16", torch.bfloat16), ("bfloat16", torch.bfloat16), ("fp16", torch.float16)):
        section = config_dict.get(key, {})
        if isinstance(section, dict) and section.get("enabled", False):
            return dt
    return None


def merge_tp_model_init_into_config(config_dict, mpu, mesh_param, dist_module):
    from ... import merge_tp_model_init_into_config as merge
    return merge(config_dict, mpu, mesh_param, dist_module)
