"""Load a universal checkpoint (``ds_to_universal`` output) into an engine of ANY dp/tp shape.

Reference: ``checkpoint/universal_checkpoint.py:149-291`` (``load_hp_checkpoint_state``: re-slice by current
tp rank/world, strip/re-pad vocab, narrow to this rank's fragment) driven by
``base_optimizer.py:332 load_hp_checkpoint_state_from_checkpoint_dir``. Here the flat-unit layout makes this a direct
copy of each parameter's overlap with the rank shard (``FlatUnit.param_fragments``).
"""
import glob
import os

import torch

from .. import comm as dist
from ..utils.logging import log_dist, logger
from .ds_to_universal import CAT_DIM, PARAM, PARAM_N_SUB_PARAMS, SUB_PARAM_SHAPE, VOCAB_TENSOR


def _tp_slice(ckpt, full_shape_local, tp_rank, tp_world):
    """Return this TP rank's slice of a merged universal tensor."""
    t = ckpt[PARAM] if isinstance(ckpt, dict) else ckpt
    if tp_world == 1 or not isinstance(ckpt, dict) or CAT_DIM not in ckpt:
        if isinstance(ckpt, dict) and ckpt.get(VOCAB_TENSOR) and t.shape[0] < full_shape_local[0]:
            pad = torch.zeros((full_shape_local[0] - t.shape[0], ) + tuple(t.shape[1:]), dtype=t.dtype)
            t = torch.cat([t, pad], 0)
        return t
    dim = ckpt[CAT_DIM]
    if ckpt.get(VOCAB_TENSOR):
        padded = full_shape_local[dim] * tp_world
        if t.shape[dim] < padded:
            shp = list(t.shape)
            shp[dim] = padded - t.shape[dim]
            t = torch.cat([t, torch.zeros(shp, dtype=t.dtype)], dim)
    n_sub = ckpt.get(PARAM_N_SUB_PARAMS)
    sub_shape = ckpt.get(SUB_PARAM_SHAPE)
    if n_sub:
        subs = t.chunk(n_sub, dim=dim)
        return torch.cat([s.chunk(tp_world, dim=dim)[tp_rank] for s in subs], dim=dim)
    if sub_shape:
        subs = t.split(list(sub_shape), dim=dim)
        return torch.cat([s.chunk(tp_world, dim=dim)[tp_rank] for s in subs], dim=dim)
    return t.chunk(tp_world, dim=dim)[tp_rank]


def load_universal_into_engine(engine, load_dir, tag, load_optimizer_states=True, load_lr_scheduler_states=True):
    folder = os.path.join(load_dir, str(tag))
    zero_dir = os.path.join(folder, "zero")
    if not os.path.isdir(zero_dir):
        raise FileNotFoundError(f"{zero_dir} is not a universal checkpoint (missing 'zero' folder)")
    z = engine.optimizer
    from ..utils import groups
    tp_world = engine.mp_world_size
    tp_rank = groups._get_model_parallel_rank() if tp_world > 1 else 0
    # module states (buffers, counters, scheduler) from any copied model_states file
    model_files = sorted(glob.glob(os.path.join(folder, "*model_states.pt")))
    client_state = {}
    if model_files:
        ms = torch.load(model_files[0], map_location="cpu", weights_only=False)
        bufs = dict(engine.module.named_buffers())
        for name, val in (ms.get("module") or {}).items():
            if name in bufs and bufs[name] is not None and torch.is_tensor(val) and val.numel() == bufs[name].numel():
                bufs[name].data.copy_(val.to(bufs[name].device))
        engine.global_steps = ms.get("global_steps", 0)
        engine.global_samples = ms.get("global_samples", 0)
        engine.skipped_steps = ms.get("skipped_steps", 0)
        if load_lr_scheduler_states and engine.lr_scheduler is not None and ms.get("lr_scheduler") is not None:
            engine.lr_scheduler.load_state_dict(ms["lr_scheduler"])
    missing = []
    with torch.no_grad():
        for u in z.units:
            for i, name in enumerate(u.names):
                pdir = os.path.join(zero_dir, name)
                if not os.path.isdir(pdir):
                    missing.append(name)
                    continue
                frags = [(s_in_p, s_in_shard, n) for (pi, s_in_p, s_in_shard, n) in u.param_fragments() if pi == i]
                if not frags:
                    continue
                for key in ("fp32", "exp_avg", "exp_avg_sq"):
                    path = os.path.join(pdir, f"{key}.pt")
                    if not os.path.exists(path) or (key != "fp32" and not load_optimizer_states):
                        continue
                    ck = torch.load(path, map_location="cpu", weights_only=False)
                    if isinstance(ck, dict) and ck.get("is_expert_param") and getattr(u, "group_key", "dp") != "dp":
                        # all experts of the layer: keep this expert-parallel rank's block (any expert-parallel degree
                        # that divides the expert count loads the same file)
                        t = ck["param"]
                        n_local = u.shapes[i][0]
                        ep_rank = groups._get_expert_parallel_rank(u.group_key) if t.shape[0] != n_local else 0
                        ck = t[ep_rank * n_local:(ep_rank + 1) * n_local]
                    full = _tp_slice(ck, u.shapes[i], tp_rank, tp_world).reshape(-1).float()
                    if full.numel() != u.numels[i]:
                        raise RuntimeError(f"universal checkpoint tensor {name}/{key} has {full.numel()} elements, "
                                           f"model expects {u.numels[i]}")
                    for (s_in_p, s_in_shard, n) in frags:
                        src = full[s_in_p:s_in_p + n]
                        if key == "fp32":
                            u.master_shard[s_in_shard:s_in_shard + n].copy_(src.to(u.master_shard.dtype))
                        else:
                            for (skey, a, b, gi, uu, v) in z._segments:
                                if uu is not u:
                                    continue
                                lo, hi = a - u.shard_offset, b - u.shard_offset
                                x, y = max(lo, s_in_shard), min(hi, s_in_shard + n)
                                if x < y:
                                    stt = z.optimizer.state[v]
                                    if key not in stt:
                                        stt[key] = torch.zeros_like(v, dtype=torch.float32)
                                    stt[key].view(-1)[x - lo:y - lo].copy_(src[x - s_in_shard:y - s_in_shard])
                step_path = os.path.join(pdir, "step.pt")
                if load_optimizer_states and os.path.exists(step_path):
                    step = torch.load(step_path, weights_only=False)
                    for (skey, a, b, gi, uu, v) in z._segments:
                        if uu is u:
                            z.optimizer.state[v]["step"] = int(step) if not torch.is_tensor(step) else int(step.item())
                            z.optimizer.param_groups[gi]["step"] = z.optimizer.state[v]["step"]
        z._lp_written_by_step = False
        z._sync_lp_from_master()
        z._post_step_gather()
    opt_path = os.path.join(zero_dir, "optimizer_state.pt")
    if load_optimizer_states and os.path.exists(opt_path):
        scal = torch.load(opt_path, map_location="cpu", weights_only=False)
        for g, saved in zip(z.optimizer.param_groups, scal.get("param_groups", [])):
            g.update({k: v for k, v in saved.items() if k != "params"})
        if "loss_scaler" in scal:
            z.loss_scaler = scal["loss_scaler"]
    if hasattr(z.optimizer, "on_state_loaded"):
        z.optimizer.on_state_loaded()
    if missing:
        logger.warning(f"universal checkpoint is missing {len(missing)} parameters, e.g. {missing[:3]}")
    log_dist(f"loaded universal checkpoint from {folder}", [0])
    return folder, client_state


# ---- reference module-level API (``checkpoint/universal_checkpoint.py``) ----------------------------
from dataclasses import dataclass  # noqa: E402
from typing import List, Tuple, Union  # noqa: E402


@dataclass
class SubparamShape:
    """Parameters matching ``patterns`` are made of fused sub-parameters: ``shape`` is the logical view whose entry at
    ``partition_dim`` lists the sub-parameter sizes (or their count); used by ``ds_to_universal`` to merge / slice the
    tensor-parallel shards of fused QKV or gate-up weights."""
    patterns: List[str]
    shape: Tuple[Union[Tuple[int], int]]
    partition_dim: int


def load_hp_checkpoint_state(self, folder, tp_rank, tp_world_size):
    """Bound onto a parameter by :func:`enable_universal_checkpoint`: read ``<folder>/fp32.pt`` (+ optimizer state
    files next to it), cut this tensor-parallel rank's slice, and write it into the parameter's master weights /
    optimizer state through the fragment API. Returns the step recorded in the files (or None)."""
    from ..utils import tensor_fragment as tf
    step = None
    for fname in sorted(os.listdir(folder)):
        if not fname.endswith(".pt"):
            continue
        key = fname[:-3]
        ckpt = torch.load(os.path.join(folder, fname), map_location="cpu", weights_only=False)
        if key == "step":
            step = ckpt
            continue
        full = _tp_slice(ckpt, tuple(getattr(self, "ds_shape", self.shape)), tp_rank, tp_world_size)
        full = full.reshape(tuple(getattr(self, "ds_shape", self.shape)))
        if key == "fp32":
            tf.safe_set_full_fp32_param(self, full)
        else:
            tf.safe_set_full_optimizer_state(self, full, key)
    return step


def enable_universal_checkpoint(param_list):
    import types
    for param in param_list:
        param.load_hp_checkpoint_state = types.MethodType(load_hp_checkpoint_state, param)
