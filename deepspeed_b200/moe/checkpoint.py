"""Per-expert checkpoint files ``layer_{L}_expert_{E}_mp_rank_{MP}_model_states.pt`` (reference
``engine.py:4243-4257`` and ``_save_moe_checkpoint``): experts are saved once per expert by the expert-DP rank 0."""
import os
import re
from collections import OrderedDict

import torch

from ..utils import groups


def _moe_layers(engine):
    from .layer import MoE
    out = []
    for name, m in engine.module.named_modules():
        if isinstance(m, MoE):
            out.append((name, m))
    return out


def _expert_prefix(name):
    return f"{name}.deepspeed_moe.experts.deepspeed_experts." if name else "deepspeed_moe.experts.deepspeed_experts."


def strip_expert_params(engine, module_sd):
    prefixes = [_expert_prefix(n) for n, _ in _moe_layers(engine)]
    return OrderedDict((k, v) for k, v in module_sd.items() if not any(k.startswith(p) for p in prefixes))


def save_moe_experts(engine, save_dir, tag):
    from .saver_names import expert_ckpt_name
    full_sd = engine.module.state_dict()
    for layer_id, (name, moe) in enumerate(_moe_layers(engine)):
        gname = moe.expert_group_name
        if groups._get_expert_data_parallel_rank(gname) != 0:
            continue
        ep_rank = groups._get_expert_parallel_rank(gname)
        prefix = _expert_prefix(name)
        for local_id in range(moe.num_local_experts):
            global_id = ep_rank * moe.num_local_experts + local_id
            sd = OrderedDict()
            for k, v in full_sd.items():
                if k.startswith(f"{prefix}{local_id}."):
                    sd[k.replace(f"{prefix}{local_id}.", f"{prefix}{global_id}.")] = v.detach().cpu().clone()
            engine.checkpoint_engine.save(sd, expert_ckpt_name(save_dir, layer_id, global_id, tag))


def load_moe_experts(engine, load_dir, tag, module_sd):
    from .saver_names import expert_ckpt_name
    for layer_id, (name, moe) in enumerate(_moe_layers(engine)):
        gname = moe.expert_group_name
        ep_rank = groups._get_expert_parallel_rank(gname)
        prefix = _expert_prefix(name)
        for local_id in range(moe.num_local_experts):
            global_id = ep_rank * moe.num_local_experts + local_id
            path = expert_ckpt_name(load_dir, layer_id, global_id, tag)
            if not os.path.exists(path):
                continue
            sd = engine.checkpoint_engine.load(path, map_location="cpu")
            for k, v in sd.items():
                module_sd[k.replace(f"{prefix}{global_id}.", f"{prefix}{local_id}.")] = v


def load_expert_files(checkpoint_path, tag, state_dict, model=None, mpu=None, num_experts=1, old_moe_load=False):
    """Static form used by ``DeepSpeedEngine.load_moe_state_dict`` (reference ``engine.py:3994``): merge the files of
    this rank's experts into ``state_dict``, renaming global expert ids to local ones. ``old_moe_load`` reads the
    layer-agnostic legacy names (``layer_id`` -1 in the reference's naming)."""
    from .layer import MoE
    from .saver_names import expert_ckpt_name
    mp_rank = 0 if mpu is None else mpu.get_model_parallel_rank()
    if old_moe_load:
        gname = groups._get_max_expert_size_name()
        ep_rank = groups._get_expert_data_parallel_rank(gname)
        n_exp = max(num_experts) if isinstance(num_experts, (list, tuple)) else int(num_experts)
        n_local = n_exp // groups._get_expert_parallel_world_size(gname)
        marker = ".deepspeed_moe.experts.deepspeed_experts."
        for local_id in range(n_local):
            global_id = ep_rank * n_local + local_id
            sd = torch.load(expert_ckpt_name(checkpoint_path, -1, global_id, tag, mp_rank), map_location="cpu",
                            weights_only=False)
            for k, v in sd.items():
                state_dict[k.replace(f"{marker}{global_id}", f"{marker}{local_id}")] = v
        return state_dict
    assert model is not None, "load_moe_state_dict needs the model to enumerate its MoE layers"
    layer_id = 0
    for name, m in model.named_modules():
        if not isinstance(m, MoE):
            continue
        gname = m.expert_group_name
        ep_rank = groups._get_expert_parallel_rank(gname)
        prefix = _expert_prefix(name)
        for local_id in range(m.num_local_experts):
            global_id = ep_rank * m.num_local_experts + local_id
            path = expert_ckpt_name(checkpoint_path, layer_id, global_id, tag, mp_rank)
            if os.path.exists(path):
                sd = torch.load(path, map_location="cpu", weights_only=False)
                for k, v in sd.items():
                    state_dict[k.replace(f"{prefix}{global_id}.", f"{prefix}{local_id}.")] = v
        layer_id += 1
    return state_dict
