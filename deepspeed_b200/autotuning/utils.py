"""Config-space helpers for the autotuner (reference ``autotuning/utils.py``)."""
import copy
import itertools
import json
import os
import re


def search_error(filename):
    """First line of a stderr log that looks like an error (OOM, traceback tail)."""
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename) as f:
        for line in f:
            for s in ("Error", "error", "ERROR"):
                if s in line:
                    return line[line.find(s):].strip()
    return None


def was_interruptted(filename):
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename) as f:
        return any("KeyboardInterrupt" in line for line in f)


def replace_dict(d, u, ignored_keys=()):
    """Recursively overwrite ``d`` with ``u``."""
    if u is not None:
        for k, v in u.items():
            if k in ignored_keys:
                continue
            if v is None:
                d.pop(k, None)
            elif isinstance(v, dict):
                d[k] = replace_dict(d.get(k, {}) if isinstance(d.get(k), dict) else {}, v, ignored_keys)
            else:
                d[k] = v
    return d


def get_list(val):
    return val if isinstance(val, list) else [val]


def _flatten(d, prefix=""):
    out = {}
    for k, v in d.items():
        key = f"{prefix}.{k}" if prefix else k
        if isinstance(v, dict):
            out.update(_flatten(v, key))
        else:
            out[key] = v
    return out


def _unflatten(flat):
    out = {}
    for key, v in flat.items():
        cur = out
        parts = key.split(".")
        for p in parts[:-1]:
            cur = cur.setdefault(p, {})
        cur[parts[-1]] = v
    return out


def get_all_configs(tuning_space, ignore_keys=None):
    """Cartesian product over every list-valued leaf of the tuning space."""
    flat = _flatten(tuning_space)
    ignore = set(ignore_keys or [])
    keys = list(flat)
    choices = [get_list(flat[k]) if k.split(".")[-1] not in ignore else [flat[k]] for k in keys]
    return [_unflatten(dict(zip(keys, combo))) for combo in itertools.product(*choices)]


def del_if_exists(t, d):
    if isinstance(d, dict):
        d.pop(t, None)
        for v in d.values():
            del_if_exists(t, v)


def canonical_name(config, tuning_keys=None, prefix="", omit_val=False):
    """Short experiment name from the tuned keys, e.g. ``z2_mbs8_gas1_rbs5e+08``."""
    flat = _flatten(config)
    parts = []
    if "zero_optimization.stage" in flat:
        parts.append(f"z{flat['zero_optimization.stage']}")
    for key, short in (("train_micro_batch_size_per_gpu", "mbs"), ("gradient_accumulation_steps", "gas")):
        if key in flat:
            parts.append(f"{short}{flat[key]}")
    for key in (tuning_keys or []):
        if key in flat and key.split(".")[-1] not in ("stage", "train_micro_batch_size_per_gpu",
                                                      "gradient_accumulation_steps"):
            short = "".join(w[0] for w in re.split(r"[_.]", key.split(".")[-1]) if w)
            v = flat[key]
            parts.append(short if omit_val else f"{short}{v:g}" if isinstance(v, float) else f"{short}{v}")
    return (prefix + "_" if prefix else "") + "_".join(parts)


def get_first_config(config):
    cfg = copy.deepcopy(config)
    for k, v in _flatten(cfg).items():
        if isinstance(v, list):
            cur = cfg
            parts = k.split(".")
            for p in parts[:-1]:
                cur = cur[p]
            cur[parts[-1]] = v[0]
    return cfg


def write_experiments(exps, exps_dir):
    os.makedirs(exps_dir, exist_ok=True)
    paths = []
    for exp in exps:
        path = os.path.join(exps_dir, f"{exp['name']}.json")
        with open(path, "w") as f:
            json.dump(exp, f)
        paths.append(path)
    return paths


def memory_to_string(n, postfix="", units=None, precision=2):
    for unit, div in (("T", 1024**4), ("G", 1024**3), ("M", 1024**2), ("K", 1024)):
        if units == unit or (units is None and n >= div):
            return f"{round(n / div, precision)} {unit}{postfix}"
    return f"{n} {postfix}"


def number_to_string(n, postfix="", units=None, precision=2):
    for unit, div in (("B", 1e9), ("M", 1e6), ("K", 1e3)):
        if units == unit or (units is None and n >= div):
            return f"{round(n / div, precision)} {unit}{postfix}"
    return f"{n} {postfix}"
