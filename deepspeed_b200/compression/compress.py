"""Entry points of the compression library (reference ``compression/compress.py``): ``init_compression`` swaps in
compression-aware layers per config group, ``redundancy_clean`` bakes the learned masks/quantization into the weights
(shrinking dimensions where a related layer consumes the pruned output), ``student_initialization`` does layer
reduction from a teacher."""
import json
import os
import re

from .config import get_compression_config
from .constants import *  # noqa: F401,F403
from .helper import (compression_preparation, fix_compression, is_module_compressible, recursive_getattr)


def check_deepspeed_config(config):
    if isinstance(config, dict):
        return config
    if isinstance(config, (str, os.PathLike)) and os.path.exists(config):
        with open(config) as f:
            return json.load(f)
    if hasattr(config, "_param_dict"):
        return config._param_dict
    raise ValueError(f"Expected a path to an existing deepspeed config or a dictionary. Received: {config}")


def get_module_name(group_name, model, key_word, exist_module_name, mpu=None, verbose=True):
    """Compressible modules whose qualified name matches the regex ``key_word`` and that no earlier group claimed."""
    found = []
    for name, module in model.named_modules():
        if re.search(key_word, name) is not None and is_module_compressible(module, mpu):
            if name in exist_module_name and verbose:
                raise ValueError(f"{name} is already added to compression, please check your config for {group_name}.")
            if name not in exist_module_name:
                exist_module_name.add(name)
                found.append(name)
    return found, exist_module_name


def get_compress_methods(model, compress_methods, mpu=None):
    """-> [[module name lists], [related module name lists], {technique: params}] per configured group."""
    out = []
    for method, content in compress_methods.items():
        if LAYER_REDUCTION in method:
            continue
        seen = set()
        shared = content[SHARED_PARAMETERS]
        for group_name, gp in content[DIFFERENT_GROUPS].items():
            names, related = [], []
            scope = gp[DIFFERENT_GROUPS_MODULE_SCOPE]
            rel_scope = gp[DIFFERENT_GROUPS_RELATED_MODULE_SCOPE]
            for i, key_word in enumerate(scope):
                found, seen = get_module_name(group_name, model, key_word, seen, mpu=mpu)
                names.append(found)
                if rel_scope:
                    related.append([get_module_name(group_name, model, rkw, set(), mpu=mpu)[0]
                                    for rkw in rel_scope[i]])
            if names:
                out.append([names, related, {method: {**gp[DIFFERENT_GROUPS_PARAMETERS], **shared}}])
    return out


def init_compression(model, deepspeed_config, teacher_model=None, mpu=None):
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    if cfg[LAYER_REDUCTION][LAYER_REDUCTION_ENABLED]:
        assert teacher_model is not None, "Teacher model is required for layer reduction"
        student_initialization(c_model, teacher_model, deepspeed_config)
    compression_preparation(c_model, get_compress_methods(c_model, cfg, mpu=mpu), mpu)
    sp = cfg[SPARSE_PRUNING][SHARED_PARAMETERS]
    if sp[SPARSE_PRUNING_ENABLED] and sp[SPARSE_PRUNING_METHOD] == SPARSE_PRUNING_METHOD_SNIP_MOMENTUM:
        from .helper import generate_pruners, register_on_step_begin
        try:
            from neural_compressor import WeightPruningConfig
        except ImportError as e:
            raise ImportError("snip_momentum sparse pruning needs the neural_compressor package") from e
        nc_cfg = WeightPruningConfig(target_sparsity=1 - sp[SPARSE_PRUNING_DENSE_RATIO],
                                     pattern=sp[SPARSE_PRUNING_BLOCK_PATTERN],
                                     pruning_frequency=sp[SPARSE_PRUNING_SCHEDULE_OFFSET_STRIDE],
                                     start_step=sp[SPARSE_PRUNING_SCHEDULE_OFFSET],
                                     end_step=sp[SPARSE_PRUNING_SCHEDULE_OFFSET_END],
                                     excluded_op_names=sp[SPARSE_PRUNING_EXCLUDED_MODULES])
        c_model.pruners = generate_pruners(nc_cfg, c_model)
        register_on_step_begin(c_model)
    return model


_CLEAN_ORDER = [WEIGHT_QUANTIZATION, SPARSE_PRUNING, ROW_PRUNING, HEAD_PRUNING, CHANNEL_PRUNING,
                ACTIVATION_QUANTIZATION]


def redundancy_clean(model, deepspeed_config, mpu=None):
    cfg = get_compression_config(check_deepspeed_config(deepspeed_config))
    c_model = model.module if hasattr(model, "module") else model
    items = sorted(get_compress_methods(c_model, cfg, mpu=mpu), key=lambda x: _CLEAN_ORDER.index(next(iter(x[2]))))
    for name_lists, related_lists, technique in items:
        need_mask = bool(related_lists)
        for i, names in enumerate(name_lists):
            masks = []
            for name in names:
                mask = fix_compression(c_model, name, technique, dim_reduction=need_mask)
                if need_mask:
                    masks.append(mask)
            if need_mask:
                for rel_names in related_lists[i]:
                    for j, name in enumerate(rel_names):
                        fix_compression(c_model, name, technique, mask=masks[j], dim_reduction=True)
    return model


def student_initialization(student_model, teacher_model, deepspeed_config):
    """Copy teacher layers ``teacher_layer[i]`` into student layer ``i`` plus the listed non-layer modules."""
    lr = get_compression_config(check_deepspeed_config(deepspeed_config))[LAYER_REDUCTION]
    prefix, teacher_layers = lr[MODULE_NAME_PREFIX], lr[TEACHER_LAYER]
    pairs = [(f"{prefix}.{s}", f"{prefix}.{t}") for s, t in enumerate(teacher_layers)]
    pairs += [(n, n) for n in lr.get(OTHER_MODULE_NAME, [])]
    for s_name, t_name in pairs:
        s_mod, t_mod = recursive_getattr(student_model, s_name), recursive_getattr(teacher_model, t_name)
        for sp, tp in zip(s_mod.parameters(), t_mod.parameters()):
            sp.data.copy_(tp.data)
