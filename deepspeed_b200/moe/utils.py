"""MoE helpers (reference ``moe/utils.py``): expert-param predicates and optimizer param-group splitting."""
from collections import defaultdict
from typing import Dict, List

import torch


def has_moe_layers(m):
    from .layer import MoE
    n = 0
    for module in m.modules():
        if isinstance(module, MoE):
            n += 1
    return n > 0, n


def is_moe_param(param):
    return hasattr(param, "allreduce") and not param.allreduce


def split_params_into_shared_and_expert_params(params):
    shared, expert = [], []
    for p in params:
        (expert if is_moe_param(p) else shared).append(p)
    return shared, expert


def split_params_grads_into_shared_and_expert_params(group):
    shared, expert = [], []
    for p in group:
        if p.grad is not None:
            (expert if is_moe_param(p) else shared).append(p.grad.to(p.dtype))
    return shared, expert


def split_params_into_different_moe_groups_for_optimizer(param_groups, max_group_size=178956971):
    """Move expert params of every group into their own groups named after the expert group
    (reference moe/utils.py:72) so ZeRO can shard them over the expert-data-parallel group."""
    if isinstance(param_groups, tuple):
        param_groups = list(param_groups)
    elif isinstance(param_groups, dict):
        param_groups = [param_groups]
    elif not isinstance(param_groups, list):
        raise ValueError(f"Unknown param group type of {type(param_groups)}")
    data_parallel_group_names = {g["name"] for g in param_groups if "name" in g}
    group_moe: Dict[str, Dict[str, dict]] = defaultdict(dict)
    for pg in param_groups:
        new_params = []
        for p in pg["params"]:
            if is_moe_param(p):
                key = p.group_name
                sub = group_moe[pg.get("name", "default")]
                if key not in sub:
                    sub[key] = {k: v for k, v in pg.items() if k not in ("name", "params")}
                    sub[key].update(name=key, moe=True, params=[])
                sub[key]["params"].append(p)
            else:
                new_params.append(p)
        pg["params"] = new_params
    for sub in group_moe.values():
        for g in sub.values():
            if max_group_size is None:
                param_groups.append(g)
                continue
            cur, cur_size, flat = [], 0, []
            for p in g["params"]:
                if cur and cur_size + p.numel() > max_group_size:
                    flat.append(cur)
                    cur, cur_size = [], 0
                cur.append(p)
                cur_size += p.numel()
            if cur:
                flat.append(cur)
            for chunk in flat:
                ng = {k: v for k, v in g.items() if k != "params"}
                ng["params"] = chunk
                param_groups.append(ng)
    return param_groups


def is_moe_param_group(param_group):
    return param_group.get("moe", False)


def configure_moe_param_groups(model_parameters):
    """Param groups an optimizer can be built from when the model has experts (reference ``moe/utils.py:155``): a plain
    list of parameters becomes one dense group that is then split into dense + per-expert-group groups; existing
    groups are split unless they already carry the ``moe`` marker."""
    assert isinstance(model_parameters, list), "model_parameters must be a list"
    for p in model_parameters:
        if not isinstance(p, (torch.Tensor, dict)):
            raise TypeError("param argument that would be given to the optimizer should be an iterable of Tensors or "
                            f"dicts, but got {type(p)}")
    first = model_parameters[0]
    if isinstance(first, torch.nn.Parameter):
        return split_params_into_different_moe_groups_for_optimizer({"params": model_parameters, "name": "dense-params"})
    if not any("moe" in g for g in model_parameters):
        return split_params_into_different_moe_groups_for_optimizer(model_parameters)
    return model_parameters
