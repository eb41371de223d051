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


def _autoep_layers(engine):
    """AutoEP blocks (stacked expert weights ``experts.w13 [E_local, 2I, H]`` / ``experts.w2 [E_local, H, I]``)."""
    from ..parallel.auto_ep import AutoEPMoELayer
    return [(name, m) for name, m in engine.module.named_modules() if isinstance(m, AutoEPMoELayer)]


_AUTOEP_KEYS = ("experts.w13", "experts.w2")


def _autoep_key(name, key):
    return f"{name}.{key}" if name else key


def strip_expert_params(engine, module_sd):
    prefixes = [_expert_prefix(n) for n, _ in _moe_layers(engine)]
    stacked = {_autoep_key(n, k) for n, _ in _autoep_layers(engine) for k in _AUTOEP_KEYS}
    return OrderedDict((k, v) for k, v in module_sd.items()
                       if k not in stacked and not any(k.startswith(p) for p in prefixes))


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
    # AutoEP blocks: one file per global expert as well, so a checkpoint loads at any expert-parallel degree. Layer
    # ids continue after the DS-MoE layers.
    base = len(_moe_layers(engine))
    for i, (name, layer) in enumerate(_autoep_layers(engine)):
        gname = layer.group_name
        if layer.ep_size > 1 and groups._get_expert_data_parallel_rank(gname) != 0:
            continue
        if layer.ep_size == 1 and getattr(engine, "seq_dp_rank", 0) != 0:
            continue
        for local_id in range(layer.num_local_experts):
            global_id = layer.ep_rank * layer.num_local_experts + local_id
            sd = OrderedDict((_autoep_key(name, k), getattr(layer.experts, k.split(".")[1])[local_id].detach().cpu().clone())
                             for k in _AUTOEP_KEYS)
            engine.checkpoint_engine.save(sd, expert_ckpt_name(save_dir, base + i, global_id, tag))


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
    base = len(_moe_layers(engine))
    for i, (name, layer) in enumerate(_autoep_layers(engine)):
        per_key = {k: [] for k in _AUTOEP_KEYS}
        for local_id in range(layer.num_local_experts):
            global_id = layer.ep_rank * layer.num_local_experts + local_id
            path = expert_ckpt_name(load_dir, base + i, global_id, tag)
            if not os.path.exists(path):
                per_key = None  # a checkpoint from before per-expert files: the stacked tensors are in module_sd
                break
            sd = engine.checkpoint_engine.load(path, map_location="cpu")
            for k in _AUTOEP_KEYS:
                per_key[k].append(sd[_autoep_key(name, k)])
        if per_key is not None:
            for k in _AUTOEP_KEYS:
                module_sd[_autoep_key(name, k)] = torch.stack(per_key[k])


def expert_optimizer_ckpt_name(ckpt_dir, tag, ep_rank, mp_rank=0):
    """Reference ``engine.py _get_optimizer_ckpt_name``: optimizer state of one expert-parallel rank when ZeRO does not
    write per-rank files."""
    return os.path.join(ckpt_dir, str(tag), f"expp_rank_{ep_rank}_mp_rank_{mp_rank:02d}_optim_states.pt")


def _max_ep(engine):
    sizes = [m.ep_size for _, m in _moe_layers(engine)] + [m.ep_size for _, m in _autoep_layers(engine)]
    return max(sizes, default=1)


def save_expert_optimizer_state(engine, save_dir, tag):
    """Without ZeRO files the optimizer state sits in rank 0's model-state file; with expert parallelism every
    expert-parallel rank owns different expert parameters, so each (first replica of it) writes its own."""
    ep = _max_ep(engine)
    if ep <= 1 or engine.optimizer is None:
        return
    gname = f"ep_size_{ep}"
    if groups._get_expert_data_parallel_rank(gname) != 0:
        return
    path = expert_optimizer_ckpt_name(save_dir, tag, groups._get_expert_parallel_rank(gname))
    engine.checkpoint_engine.save({"optimizer": engine.optimizer.state_dict()}, path)


def load_expert_optimizer_state(engine, load_dir, tag):
    """This rank's ``expp_rank_*`` optimizer state, or ``None``."""
    ep = _max_ep(engine)
    if ep <= 1:
        return None
    path = expert_optimizer_ckpt_name(load_dir, tag, groups._get_expert_parallel_rank(f"ep_size_{ep}"))
    if not os.path.exists(path):
        return None
    return engine.checkpoint_engine.load(path, map_location="cpu")["optimizer"]


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
