#!/usr/bin/env python
"""Consolidate a ZeRO-1/2/3 checkpoint into a single fp32 ``state_dict`` (offline, CPU only, no deepspeed import).

This script is copied into every checkpoint folder. It understands the on-disk layout produced by this framework
and by the reference (same files/keys, see ``utils/zero_to_fp32.py:282-460`` of the reference):

* ``*_optim_states.pt`` per DP rank with ``optimizer_state_dict`` -> ``zero_stage``, ``partition_count`` and
  ``single_partition_of_fp32_groups`` (stage 1/2: per param-group contiguous rank slices of the flat fp32 vector,
  padded to ``2*world``) or ``fp32_flat_groups`` (stage 3: every parameter split ``ceil(numel/world)`` per rank);
* ``*_model_states.pt`` with ``param_shapes``, ``buffer_names``, ``shared_params``, frozen-parameter info.

Usage::

    python zero_to_fp32.py <checkpoint_dir> <output_dir_or_file> [-t TAG] [--safe_serialization] [--max_shard_size 5GB]
"""
import argparse
import glob
import math
import os
import re
from collections import OrderedDict

import torch

ZERO_STAGE = "zero_stage"
PARTITION_COUNT = "partition_count"
OPTIMIZER_STATE_DICT = "optimizer_state_dict"
SINGLE_PARTITION_OF_FP32_GROUPS = "single_partition_of_fp32_groups"
FP32_FLAT_GROUPS = "fp32_flat_groups"
PARAM_SHAPES = "param_shapes"
BUFFER_NAMES = "buffer_names"
FROZEN_PARAM_SHAPES = "frozen_param_shapes"
FROZEN_PARAM_FRAGMENTS = "frozen_param_fragments"


def _natural_key(text):
    return [int(c) if c.isdigit() else c for c in re.split(r"(\d+)", text)]


def _files(ckpt_dir, pattern):
    out = sorted(glob.glob(os.path.join(ckpt_dir, pattern)), key=_natural_key)
    if not out:
        raise FileNotFoundError(f"can't find {pattern} files in directory '{ckpt_dir}'")
    return out


class _Stub:
    """Placeholder for pickled helper objects (loss scaler, fragment addresses) whose class is not importable."""

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        self.__dict__.update(state if isinstance(state, dict) else {})


class _TolerantUnpickler:
    """pickle module shim for torch.load: unknown classes become :class:`_Stub` so the script runs stand-alone."""
    import pickle as _p
    __name__ = "tolerant_pickle"

    class Unpickler(_p.Unpickler):

        def find_class(self, mod_name, name):
            try:
                return super().find_class(mod_name, name)
            except (ImportError, AttributeError):
                return _Stub

    load = staticmethod(_p.load)
    loads = staticmethod(_p.loads)


def _load(path):
    try:
        return torch.load(path, map_location="cpu", weights_only=False)
    except (ImportError, AttributeError):
        return torch.load(path, map_location="cpu", weights_only=False, pickle_module=_TolerantUnpickler)


def _resolve_tag(checkpoint_dir, tag):
    if tag is None:
        latest = os.path.join(checkpoint_dir, "latest")
        if not os.path.isfile(latest):
            raise ValueError(f"Unable to find 'latest' file at {latest}")
        with open(latest) as fd:
            tag = fd.read().strip()
    ds_dir = os.path.join(checkpoint_dir, tag)
    if not os.path.isdir(ds_dir):
        raise FileNotFoundError(f"Directory '{ds_dir}' doesn't exist")
    return ds_dir


def _cdiv(a, b):
    return (a + b - 1) // b


def _merge_stage12(fp32_groups_per_rank, param_shapes, world):
    """fp32_groups_per_rank[rank][group] -> {name: tensor}."""
    out = OrderedDict()
    n_groups = len(fp32_groups_per_rank[0])
    for g in range(n_groups):
        full = torch.cat([fp32_groups_per_rank[r][g].reshape(-1) for r in range(world)], 0)
        off = 0
        for name, shape in param_shapes[g].items():
            n = int(torch.Size(shape).numel())
            out[name] = full.narrow(0, off, n).view(shape).clone()
            off += n
        # sanity: what is left must be alignment padding only
        align = 2 * world
        if _cdiv(off, align) * align < full.numel() - align:
            raise ValueError(f"group {g}: consumed {off} of {full.numel()} elements - layout mismatch")
    return out


def zero3_partitioned_param_info(unpartitioned_numel, world_size):
    """(elements per rank, padding added to the last rank) of a ZeRO-3 parameter."""
    remainder = unpartitioned_numel % world_size
    return _cdiv(unpartitioned_numel, world_size), (world_size - remainder) if remainder else 0


class GatheredTensor:
    """A parameter of a ZeRO-3 checkpoint that is put together only when asked for (reference
    ``utils/zero_to_fp32.py:421``): it remembers where its pieces sit in the ranks' flat buffers. With ``lazy_mode`` the
    state dict holds these instead of tensors, so a conversion keeps one parameter in memory at a time."""

    def __init__(self, flat_per_rank, offset, per_rank_numel, shape):
        self.flat_per_rank, self.offset, self.per, self.shape = flat_per_rank, offset, per_rank_numel, torch.Size(shape)
        self.dtype = flat_per_rank[0].dtype

    def numel(self):
        return self.shape.numel()

    def contiguous(self):
        parts = [f.narrow(0, self.offset, self.per) for f in self.flat_per_rank]
        return torch.cat(parts, 0).narrow(0, 0, self.numel()).view(self.shape).clone()


def to_torch_tensor(state_dict, return_empty_tensor=False):
    """Materialise the :class:`GatheredTensor` entries of a lazily built state dict (``return_empty_tensor``: only
    shapes and dtypes, for planning the shards)."""
    out = OrderedDict()
    seen = {}
    for name, t in state_dict.items():
        if isinstance(t, GatheredTensor):
            if id(t) in seen:
                out[name] = out[seen[id(t)]]  # shared parameters stay shared
                continue
            seen[id(t)] = name
            out[name] = torch.empty(t.shape, dtype=t.dtype) if return_empty_tensor else t.contiguous()
        else:
            out[name] = t
    return out


def _merge_stage3(flat_per_rank, param_shapes, world, lazy=False):
    """flat_per_rank[rank] = concatenation of all sub-group flats of that rank."""
    out = OrderedDict()
    shapes = OrderedDict()
    for d in param_shapes:
        shapes.update(d)
    off = 0
    for name, shape in shapes.items():
        n = int(torch.Size(shape).numel())
        per, _ = zero3_partitioned_param_info(n, world)
        g = GatheredTensor(flat_per_rank, off, per, shape)
        out[name] = g if lazy else g.contiguous()
        off += per
    return out


def _merge_expert_units(optim_states, key="fp32"):
    """Expert parameters are not in the data-parallel flats: every rank file carries its expert units (this rank's shard
    of the unit's fp32 master, plus moments). Shards of a unit are concatenated over its expert-data-parallel ranks,
    cut into parameters, and the expert-parallel ranks' tensors are put together: stacked experts (``[E_local, ...]``)
    along dim 0, per-expert module names by renumbering the local expert index to the global one."""
    import re
    shards = {}  # (unit name, ep_rank) -> {edp rank: entry}
    for o in optim_states:
        for uname, e in (o.get("expert_units") or {}).items():
            shards.setdefault((uname, e.get("ep_rank", 0)), {})[e["rank"]] = e
    per_rank = {}  # param name -> {ep_rank: tensor}
    ep_size = 1
    for (uname, ep_rank), by_rank in shards.items():
        first = by_rank[min(by_rank)]
        if len(by_rank) != first["world"] or "shapes" not in first:
            continue  # incomplete set of shards (or an old file without geometry): leave these parameters out
        ep_size = max(ep_size, first.get("ep_size", 1))
        pick = (lambda e: e["fp32"]) if key == "fp32" else (lambda e: e["state"][key])
        if key != "fp32" and any(key not in e["state"] for e in by_rank.values()):
            continue
        flat = torch.cat([pick(by_rank[r]).float().reshape(-1) for r in sorted(by_rank)])
        for name, shape, off in zip(first["names"], first["shapes"], first["offsets"]):
            n = torch.Size(shape).numel()
            per_rank.setdefault(name, {})[ep_rank] = flat[off:off + n].view(shape).clone()
    out = {}
    for name, by_ep in per_rank.items():
        ranks = sorted(by_ep)
        m = re.search(r"deepspeed_experts\.(\d+)\.", name)
        if m is None:  # stacked experts: local slices along dim 0
            out[name] = torch.cat([by_ep[r] for r in ranks], dim=0) if len(ranks) > 1 else by_ep[ranks[0]]
        else:
            prefix = name[:m.start()]
            n_local = 1 + max(int(mm.group(1)) for nm in per_rank
                              for mm in [re.search(r"deepspeed_experts\.(\d+)\.", nm)]
                              if mm is not None and nm.startswith(prefix))
            for r in ranks:
                gid = r * n_local + int(m.group(1))
                out[name[:m.start(1)] + str(gid) + name[m.end(1):]] = by_ep[r]
    return out


def get_checkpoint_files(checkpoint_dir, glob_pattern):
    files = _files(checkpoint_dir, glob_pattern)
    if not files:
        raise FileNotFoundError(f"can't find {glob_pattern} files in directory '{checkpoint_dir}'")
    return files


def get_optim_files(checkpoint_dir):
    return get_checkpoint_files(checkpoint_dir, "*_optim_states.pt")


def get_model_state_files(checkpoint_dir):
    return [f for f in get_checkpoint_files(checkpoint_dir, "*_model_states.pt")
            if "_expert_" not in os.path.basename(f) and not os.path.basename(f).startswith("expert_")]


def get_model_state_file(checkpoint_dir, zero_stage):
    if not os.path.isdir(checkpoint_dir):
        raise FileNotFoundError(f"Directory '{checkpoint_dir}' doesn't exist")
    name = "mp_rank_00_model_states.pt" if zero_stage <= 2 else "zero_pp_rank_0_mp_rank_00_model_states.pt"
    path = os.path.join(checkpoint_dir, name)
    if not os.path.exists(path):
        raise FileNotFoundError(f"can't find model states file at '{path}'")
    return path


def get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag=None, exclude_frozen_parameters=False,
                                             lazy_mode=False):
    """Return the consolidated fp32 state dict (parameters + buffers + shared-parameter aliases)."""
    ds_dir = _resolve_tag(checkpoint_dir, tag)
    optim_files = _files(ds_dir, "*_optim_states.pt")
    model_files = [f for f in _files(ds_dir, "*_model_states.pt")
                   if "_expert_" not in os.path.basename(f) and not os.path.basename(f).startswith("expert_")]
    optim = [_load(f)[OPTIMIZER_STATE_DICT] for f in optim_files]
    stage = optim[0][ZERO_STAGE]
    world = optim[0][PARTITION_COUNT]
    world = max(world) if isinstance(world, (list, tuple)) else world
    if world != len(optim_files):
        raise ValueError(f"Expected {world} of '*_optim_states.pt' under '{ds_dir}' but found {len(optim_files)} files.")
    model_state = _load(model_files[0])
    if BUFFER_NAMES not in model_state:
        raise ValueError(f"{model_files[0]} is not a model state checkpoint")
    param_shapes = model_state[PARAM_SHAPES]
    print(f"Detected checkpoint of type zero stage {stage}, world_size: {world}")
    if stage <= 2:
        groups = [[t.float() for t in o[SINGLE_PARTITION_OF_FP32_GROUPS]] for o in optim]
        state = _merge_stage12(groups, param_shapes, world)
    elif stage == 3:
        flats = [torch.cat([t.float().reshape(-1) for t in o[FP32_FLAT_GROUPS]], 0) for o in optim]
        state = _merge_stage3(flats, param_shapes, world, lazy=lazy_mode)
    else:
        raise ValueError(f"unknown zero stage {stage}")
    state.update(_merge_expert_units(optim))
    # buffers (restored to fp32), frozen params, shared params
    buffers = {k: v.float() for k, v in model_state["module"].items() if k in set(model_state[BUFFER_NAMES])}
    state.update(buffers)
    if not exclude_frozen_parameters and model_state.get(FROZEN_PARAM_SHAPES):
        frags = model_state.get(FROZEN_PARAM_FRAGMENTS) or {}
        for name, shape in model_state[FROZEN_PARAM_SHAPES].items():
            if stage == 3 and len(model_files) > 1:
                pieces = [(_load(f).get(FROZEN_PARAM_FRAGMENTS) or {}).get(name) for f in model_files]
                if all(p is not None for p in pieces) and sum(p.numel() for p in pieces) >= torch.Size(shape).numel():
                    state[name] = torch.cat([p.reshape(-1) for p in pieces])[:torch.Size(shape).numel()].view(shape).float()
                    continue
            if name in frags and frags[name].numel() == torch.Size(shape).numel():
                state[name] = frags[name].view(shape).float()
    for alias, src in (model_state.get("shared_params") or {}).items():
        if src in state:
            state[alias] = state[src]
    return state


def _parse_size(s):
    if isinstance(s, int):
        return s
    m = re.match(r"^\s*(\d+(?:\.\d+)?)\s*([KMGT]?B)\s*$", s.upper())
    if not m:
        raise ValueError(f"bad size {s}")
    return int(float(m.group(1)) * {"B": 1, "KB": 2**10, "MB": 2**20, "GB": 2**30, "TB": 2**40}[m.group(2)])


def convert_zero_checkpoint_to_fp32_state_dict(checkpoint_dir, output_dir, max_shard_size="5GB",
                                               safe_serialization=False, tag=None, exclude_frozen_parameters=False):
    """Write the consolidated weights as (sharded) ``pytorch_model*.bin`` or ``model*.safetensors`` + index json."""
    # lazily gathered (ZeRO-3): a parameter is put together when its shard is written, one shard in memory at a time
    state = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag, exclude_frozen_parameters, lazy_mode=True)
    if output_dir.endswith((".bin", ".pt")):
        os.makedirs(os.path.dirname(os.path.abspath(output_dir)), exist_ok=True)
        torch.save(to_torch_tensor(state), output_dir)
        return
    os.makedirs(output_dir, exist_ok=True)
    limit = _parse_size(max_shard_size)
    shards, cur, cur_bytes = [], OrderedDict(), 0
    for k, v in state.items():
        nb = v.numel() * (4 if isinstance(v, GatheredTensor) else v.element_size())
        if cur and cur_bytes + nb > limit:
            shards.append(cur)
            cur, cur_bytes = OrderedDict(), 0
        cur[k] = v
        cur_bytes += nb
    if cur:
        shards.append(cur)
    base, ext = ("model", ".safetensors") if safe_serialization else ("pytorch_model", ".bin")
    index = {"metadata": {"total_size": sum(v.numel() * (4 if isinstance(v, GatheredTensor) else v.element_size())
                                           for v in state.values())}, "weight_map": {}}
    for i, sh in enumerate(shards):
        sh = to_torch_tensor(sh)
        name = f"{base}{ext}" if len(shards) == 1 else f"{base}-{i + 1:05d}-of-{len(shards):05d}{ext}"
        path = os.path.join(output_dir, name)
        if safe_serialization:
            from safetensors.torch import save_file
            save_file({k: v.contiguous() for k, v in sh.items()}, path, metadata={"format": "pt"})
        else:
            torch.save(sh, path)
        for k in sh:
            index["weight_map"][k] = name
    if len(shards) > 1:
        import json
        with open(os.path.join(output_dir, f"{base}{ext}.index.json"), "w") as f:
            json.dump(index, f, indent=2, sort_keys=True)


def load_state_dict_from_zero_checkpoint(model, checkpoint_dir, tag=None):
    """Load consolidated fp32 weights into ``model`` (on CPU)."""
    state = get_fp32_state_dict_from_zero_checkpoint(checkpoint_dir, tag)
    model = model.cpu()
    model.load_state_dict(state, strict=False)
    return model


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("checkpoint_dir", type=str, help="path to the checkpoint folder (contains 'latest')")
    parser.add_argument("output_dir", type=str, help="output directory (or .bin/.pt file) for the fp32 weights")
    parser.add_argument("--max_shard_size", type=str, default="5GB")
    parser.add_argument("--safe_serialization", default=False, action="store_true")
    parser.add_argument("-t", "--tag", type=str, default=None)
    parser.add_argument("--exclude_frozen_parameters", action="store_true")
    parser.add_argument("-d", "--debug", action="store_true")
    a = parser.parse_args()
    convert_zero_checkpoint_to_fp32_state_dict(a.checkpoint_dir, a.output_dir, max_shard_size=a.max_shard_size,
                                               safe_serialization=a.safe_serialization, tag=a.tag,
                                               exclude_frozen_parameters=a.exclude_frozen_parameters)
