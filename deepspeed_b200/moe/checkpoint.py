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
