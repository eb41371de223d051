#!/usr/bin/env python
"""Convert a ZeRO checkpoint (any DP/TP/PP degree) into the *universal* layout: one folder per parameter.

Output (same as the reference ``checkpoint/ds_to_universal.py`` ``main`` :1017)::

    <out>/zero/<param_name>/fp32.pt, exp_avg.pt, exp_avg_sq.pt, step.pt     dict {"param": tensor, "cat_dim": ...}
    <out>/zero/optimizer_state.pt                                            non-sharded optimizer scalars + groups
    <out>/mp_rank_*_model_states.pt (copied)                                 module/buffers/metadata
    <root>/latest_universal                                                  text file = output folder name

Pipeline: (1) *extract* every rank's flat partition into per-parameter fragments using ``param_slice_mappings``
(stage 1/2) or ``param_shapes`` + ``ceil(numel/world)`` (stage 3); (2) *merge* DP fragments by concatenation and TP
slices by the pattern lists in ``universal_checkpoint_info`` (replicated / averaged / row-parallel cat dim 1 /
default cat dim 0 / sub-params / vocab padding).
"""
import argparse
import glob
import os
import re
import shutil
from collections import OrderedDict
from concurrent.futures import ProcessPoolExecutor

import torch

UNIVERSAL_CHECKPOINT_INFO = "universal_checkpoint_info"
PARAM = "param"
EP_IS_EXPERT_PARAM = "is_expert_param"
EP_NUM_EXPERTS = "ep_num_experts"
CAT_DIM = "cat_dim"
VOCAB_TENSOR = "vocab_tensor"
PARAM_N_SUB_PARAMS = "param_n_sub_params"
SUB_PARAM_SHAPE = "sub_param_shape"
VOCABULARY_PARAMETER_PATTERNS = "vocabulary_parameter_patterns"
PIPELINE_REPLICATED_PARAMETER_PATTERNS = "pipeline_replicated_parameter_patterns"
TP_REPLICATED_PARAMETER_PATTERNS = "tp_replicated_parameter_patterns"
PARAMETER_TO_AVERAGE_PATTERNS = "parameter_to_average_patterns"
PARAMETER_WITH_ROW_PARALLELISM_PATTERNS = "parameter_with_row_parallelism_patterns"
PARAMETER_WITH_2_SUB_PARAMS_CAT_DIM_0 = "parameter_with_2_sub_params_cat_dim_0"
PARAMETER_WITH_SUB_PARAMS = "parameter_with_sub_params"
ORIGINAL_VOCAB_SIZE = "original_vocab_size"
PADDED_VOCAB_SIZE = "padded_vocab_size"
STATE_KEYS = ("fp32", "exp_avg", "exp_avg_sq")


def parse_arguments(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--input_folder", type=str, required=True, help="Input DeepSpeed Checkpoint folder (tag dir)")
    p.add_argument("--output_folder", type=str, required=True, help="Output universal checkpoint folder")
    p.add_argument("--num_extract_workers", default=4, type=int)
    p.add_argument("--num_merge_workers", default=2, type=int)
    p.add_argument("--keep_temp_folder", action="store_true")
    p.add_argument("--no_strict", dest="strict", action="store_false")
    p.add_argument("--inject_missing_state", action="store_true",
                   help="create zero optimizer state for parameters missing it (e.g. frozen during training)")
    return p.parse_args(argv)


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False)


def _natural(text):
    return [int(c) if c.isdigit() else c for c in re.split(r"(\d+)", text)]


def _cdiv(a, b):
    return (a + b - 1) // b


def _rank_files(folder):
    """{tp_rank: [optim file per dp rank]} and {tp_rank: model state file}."""
    optim = sorted(glob.glob(os.path.join(folder, "*zero_pp_rank_*_mp_rank_*_optim_states.pt")), key=_natural)
    by_tp = OrderedDict()
    for f in optim:
        m = re.search(r"zero_pp_rank_(\d+)_mp_rank_(\d+)_optim_states", os.path.basename(f))
        by_tp.setdefault(int(m.group(2)), {})[int(m.group(1))] = f
    model = {}
    for f in sorted(glob.glob(os.path.join(folder, "*mp_rank_*_model_states.pt"))):
        if "_expert_" in os.path.basename(f) or os.path.basename(f).startswith("expert_"):
            continue  # per-expert weight files of MoE layers
        m = re.search(r"mp_rank_(\d+)_model_states", os.path.basename(f))
        model.setdefault(int(m.group(1)), f)
    return {tp: [d[k] for k in sorted(d)] for tp, d in by_tp.items()}, model


def _extract_one(args):
    """Slice one rank's flat partitions into per-parameter fragment files under tmp/<name>/<tp>/<state>.<dp>."""
    tmp, tp, dp, path, model_path = args
    sd = _load(path)["optimizer_state_dict"]
    stage = sd["zero_stage"]
    world = sd["partition_count"]
    world = max(world) if isinstance(world, (list, tuple)) else world
    if stage <= 2:
        fp32 = sd["single_partition_of_fp32_groups"]
        states = sd["base_optimizer_state"]["state"]
        maps = sd["param_slice_mappings"]
        for gi, mapping in enumerate(maps):
            flat = {"fp32": fp32[gi].float()}
            for k in ("exp_avg", "exp_avg_sq"):
                if gi in states and k in states[gi]:
                    flat[k] = states[gi][k].float()
            for name, frag in mapping.items():
                start, numel = frag.start, frag.numel
                for k, t in flat.items():
                    _dump(tmp, name, tp, k, dp, t.narrow(0, start, numel).clone())
                step = states.get(gi, {}).get("step")
                if step is not None:
                    _dump(tmp, name, tp, "step", dp, torch.tensor(float(step)) if not torch.is_tensor(step) else step)
    else:
        shapes = OrderedDict()
        for d in _load(model_path)["param_shapes"]:
            shapes.update(d)
        flat = {"fp32": torch.cat([t.float().reshape(-1) for t in sd["fp32_flat_groups"]])}
        ostate = sd["optimizer_state_dict"]["state"]
        keys = sorted(ostate.keys())
        for k in ("exp_avg", "exp_avg_sq"):
            if keys and k in ostate[keys[0]]:
                flat[k] = torch.cat([ostate[i][k].float().reshape(-1) for i in keys])
        step = ostate[keys[0]].get("step") if keys else None
        off = 0
        for name, shape in shapes.items():
            n = int(torch.Size(shape).numel())
            per = _cdiv(n, world)
            for k, t in flat.items():
                _dump(tmp, name, tp, k, dp, t.narrow(0, off, per).clone())
            if step is not None:
                _dump(tmp, name, tp, "step", dp, torch.tensor(float(step)) if not torch.is_tensor(step) else step)
            off += per
    return path


def _dump(tmp, name, tp, state, dp, tensor):
    d = os.path.join(tmp, name, str(tp))
    os.makedirs(d, exist_ok=True)
    torch.save(tensor, os.path.join(d, f"{state}.{dp:02d}"))


def _merge_dp(tmp, name, tp, state, shape_numel=None):
    files = sorted(glob.glob(os.path.join(tmp, name, str(tp), f"{state}.*")), key=_natural)
    if not files:
        return None
    if state == "step":
        return torch.load(files[0], weights_only=False)
    flat = torch.cat([torch.load(f, weights_only=False).reshape(-1) for f in files])
    return flat[:shape_numel] if shape_numel is not None else flat


def _matches(patterns, name):
    return any(re.match(p, name) for p in (patterns or []))


def _merge_param(args):
    tmp, out_zero, name, shapes_by_tp, uc_info, tp_degree = args
    os.makedirs(os.path.join(out_zero, name), exist_ok=True)
    for state in STATE_KEYS + ("step", ):
        slices = []
        for tp in range(tp_degree):
            shape = shapes_by_tp[tp]
            t = _merge_dp(tmp, name, tp, state, int(torch.Size(shape).numel()) if state != "step" else None)
            if t is None:
                break
            slices.append(t if state == "step" else t.view(shape))
        if len(slices) != tp_degree:
            continue
        if state == "step":
            torch.save(slices[0], os.path.join(out_zero, name, "step.pt"))
            continue
        ckpt = {}
        if tp_degree == 1 or _matches(uc_info.get(TP_REPLICATED_PARAMETER_PATTERNS), name):
            param = slices[0]
        elif _matches(uc_info.get(PARAMETER_TO_AVERAGE_PATTERNS), name):
            param = sum(slices) / len(slices)
        elif _matches(uc_info.get(PARAMETER_WITH_2_SUB_PARAMS_CAT_DIM_0), name):
            halves = [s.chunk(2, dim=0) for s in slices]
            param = torch.cat([torch.cat([h[i] for h in halves], dim=0) for i in range(2)], dim=0)
            ckpt[CAT_DIM], ckpt[PARAM_N_SUB_PARAMS] = 0, 2
        else:
            cat_dim = 1 if _matches(uc_info.get(PARAMETER_WITH_ROW_PARALLELISM_PATTERNS), name) else 0
            sub = None
            for entry in uc_info.get(PARAMETER_WITH_SUB_PARAMS, []) or []:
                if _matches(entry.get("patterns"), name):
                    sub = entry
            if sub is not None:
                cat_dim = sub.get("partition_dim", 0)
                sizes = sub["shape"]  # full sub-param sizes along partition_dim
                per_rank = [[sz // tp_degree for sz in sizes]] * tp_degree
                pieces = [s.split(per_rank[r], dim=cat_dim) for r, s in enumerate(slices)]
                param = torch.cat([torch.cat([p[i] for p in pieces], dim=cat_dim) for i in range(len(sizes))],
                                  dim=cat_dim)
                ckpt[SUB_PARAM_SHAPE] = sizes
            else:
                param = torch.cat(slices, dim=cat_dim)
            ckpt[CAT_DIM] = cat_dim
        if _matches(uc_info.get(VOCABULARY_PARAMETER_PATTERNS), name):
            orig = uc_info.get(ORIGINAL_VOCAB_SIZE)
            if orig is not None and param.shape[0] > orig:
                param = param[:orig]
            ckpt[VOCAB_TENSOR] = True
        ckpt[PARAM] = param.clone()
        torch.save(ckpt, os.path.join(out_zero, name, f"{state}.pt"))
    return name


def main(args=None):
    args = args or parse_arguments()
    inp, out = args.input_folder.rstrip("/"), args.output_folder.rstrip("/")
    print(f"Convert DeepSpeed Checkpoint to Universal Checkpoint: {inp} -> {out}")
    optim_by_tp, model_by_tp = _rank_files(inp)
    if not optim_by_tp:
        raise FileNotFoundError(f"no *_optim_states.pt files under {inp}")
    tp_degree = len(optim_by_tp)
    tmp = os.path.join(out, "tmp")
    out_zero = os.path.join(out, "zero")
    os.makedirs(tmp, exist_ok=True)
    os.makedirs(out_zero, exist_ok=True)
    first_model = _load(model_by_tp[sorted(model_by_tp)[0]]) if model_by_tp else _load(
        sorted(glob.glob(os.path.join(inp, "*model_states.pt")), key=_natural)[0])
    uc_info = first_model.get(UNIVERSAL_CHECKPOINT_INFO, {}) or {}
    # ---- 1. extract ----
    jobs = []
    for tp, files in optim_by_tp.items():
        zero3_models = sorted(glob.glob(os.path.join(inp, f"zero_pp_rank_*_mp_rank_{tp:02d}_model_states.pt")), key=_natural)
        mpath = model_by_tp.get(tp) or (zero3_models[0] if zero3_models else None)
        for dp, f in enumerate(files):
            jobs.append((tmp, tp, dp, f, mpath))
    print(f"*** 1. Extracting ZeRO fragments ({len(jobs)} shard files)")
    if args.num_extract_workers > 1 and len(jobs) > 1:
        with ProcessPoolExecutor(max_workers=args.num_extract_workers) as ex:
            list(ex.map(_extract_one, jobs))
    else:
        for j in jobs:
            _extract_one(j)
    # ---- 2. merge ----
    shapes_by_tp = {}
    for tp in optim_by_tp:
        zero3_models = sorted(glob.glob(os.path.join(inp, f"zero_pp_rank_*_mp_rank_{tp:02d}_model_states.pt")), key=_natural)
        ms = _load(model_by_tp[tp]) if tp in model_by_tp else _load(zero3_models[0])
        d = OrderedDict()
        for g in ms["param_shapes"]:
            d.update(g)
        shapes_by_tp[tp] = d
    names = list(shapes_by_tp[sorted(shapes_by_tp)[0]].keys())
    print(f"*** 2. Merging slices of {len(names)} parameters (tp degree {tp_degree})")
    mjobs = [(tmp, out_zero, n, {tp: shapes_by_tp[tp][n] for tp in shapes_by_tp}, uc_info, tp_degree) for n in names]
    if args.num_merge_workers > 1 and len(mjobs) > 1:
        with ProcessPoolExecutor(max_workers=args.num_merge_workers) as ex:
            list(ex.map(_merge_param, mjobs))
    else:
        for j in mjobs:
            _merge_param(j)
    # ---- 2b. expert-parallel parameters: every rank file carries its expert units; put the experts of all
    # expert-parallel ranks together so the result loads at any expert-parallel degree (reference autoep_universal.py)
    from .zero_to_fp32 import _merge_expert_units
    all_optim = [_load(f)["optimizer_state_dict"] for tp in sorted(optim_by_tp) for f in optim_by_tp[tp]]
    n_expert = 0
    for state in ("fp32", "exp_avg", "exp_avg_sq"):
        for name, tensor in _merge_expert_units(all_optim, key=state).items():
            os.makedirs(os.path.join(out_zero, name), exist_ok=True)
            torch.save({PARAM: tensor.clone(), EP_IS_EXPERT_PARAM: True, EP_NUM_EXPERTS: int(tensor.shape[0])},
                       os.path.join(out_zero, name, f"{state}.pt"))
            if state == "fp32":
                n_expert += 1
                base0 = all_optim[0].get("base_optimizer_state") or all_optim[0].get("optimizer_state_dict") or {}
                step = next((st["step"] for st in (base0.get("state") or {}).values() if "step" in st), None)
                if step is not None:
                    torch.save(step, os.path.join(out_zero, name, "step.pt"))
    if n_expert:
        print(f"*** 2b. Consolidated {n_expert} expert-parallel parameters")
    # ---- 3. optimizer scalars, model states, latest_universal ----
    osd = _load(optim_by_tp[sorted(optim_by_tp)[0]][0])["optimizer_state_dict"]
    base = osd.get("base_optimizer_state") or osd.get("optimizer_state_dict") or {}
    scalars = {k: v for k, v in osd.items() if not isinstance(v, (list, dict)) or k in ("loss_scaler", )}
    scalars["param_groups"] = base.get("param_groups", [])
    torch.save(scalars, os.path.join(out_zero, "optimizer_state.pt"))
    for f in glob.glob(os.path.join(inp, "*model_states.pt")):
        shutil.copy2(f, os.path.join(out, os.path.basename(f)))
    if not args.keep_temp_folder:
        shutil.rmtree(tmp, ignore_errors=True)
    root = os.path.dirname(out)
    with open(os.path.join(root, "latest_universal"), "w") as f:
        f.write(os.path.basename(out))
    print("*** Done!")


if __name__ == "__main__":
    main()
