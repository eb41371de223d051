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


# ---- further helpers of the reference's ``autotuning/utils.py`` ---------------------------------------------
import json as _json  # noqa: E402
import re as _re  # noqa: E402

was_interrupted = was_interruptted


def find_replace_str(value, replace_dict):
    """Substitute ``$VAR`` placeholders (``$NVME_PATH`` is left for the launcher). A placeholder bound to a non-string
    must be the whole value: the value then becomes that object."""
    if not isinstance(value, str):
        return str(value)
    found = _re.findall(r"\$[\w]+", value)
    for var in found:
        key = var[1:].lower()
        if key == "nvme_path":
            continue
        assert key in replace_dict, f"unknown var key: {key}, in {replace_dict}"
        if isinstance(replace_dict[key], str):
            value = value.replace(var, replace_dict[key])
        else:
            assert len(found) == 1, "unable to replace multiple non-string matches"
            value = replace_dict[key]
    return value


def find_replace(target, replace_dict):
    """:func:`find_replace_str` over every string of a nested dict / list, in place."""
    if isinstance(target, dict):
        for k, v in target.items():
            if isinstance(v, str):
                target[k] = find_replace_str(v, replace_dict)
            elif isinstance(v, list):
                v[:] = [find_replace_str(x, replace_dict) for x in v]
            elif isinstance(v, dict):
                find_replace(v, replace_dict)
    elif isinstance(target, list):
        target[:] = [str(find_replace_str(x, replace_dict)) for x in target]


def combine_dict(d, u):
    """Merge ``u`` into ``d``: nested dicts recursively, equal keys collect their values in a list (no duplicates)."""
    for k, v in u.items():
        if isinstance(v, dict):
            d[k] = combine_dict(d.get(k, {}), v)
        elif k not in d:
            d[k] = v
        else:
            if not isinstance(d[k], list):
                d[k] = [d[k]]
            d[k].extend(x for x in get_list(v) if x not in d[k])
    return d


def get_val_by_key(d, k):
    """First value stored under ``k`` anywhere in the nested dict."""
    if k in d:
        return d[k]
    for v in d.values():
        if isinstance(v, dict):
            r = get_val_by_key(v, k)
            if r is not None:
                return r
    return None


def set_val_by_key(d, k, vv):
    """Set every occurrence of key ``k`` in the nested dict."""
    if k in d:
        d[k] = vv
    for v in d.values():
        if isinstance(v, dict):
            set_val_by_key(v, k, vv)


def validate_ds_config(config):
    """Configurations the tuner should not try: ZeRO-2 with both legacy offload flags, offload without an optimizer."""
    zero = config.get("zero_optimization") or {}
    stage = zero.get("stage")
    if stage == 2 and zero.get("cpu_offload") and zero.get("cpu_offload_params"):
        return False
    if stage == 3:
        offload = any((zero.get(k) or {}).get("device") in ("cpu", "nvme") for k in ("offload_optimizer",
                                                                                     "offload_param"))
        if offload and not config.get("optimizer"):
            return False
    return True


def remove_dupe_dicts(items):
    """Unique (nested) dictionaries of a list; order of first appearance is kept."""
    seen, out = set(), []
    for d in items:
        key = _json.dumps(d, sort_keys=True)
        if key not in seen:
            seen.add(key)
            out.append(_json.loads(key))
    return out


def prune_config(config, ignored_keys=()):
    """Delete the sections named in ``ignored_keys`` wherever they first occur on a path, in place."""
    def drop(d, k):
        if k in d:
            del d[k]
        else:
            for v in d.values():
                if isinstance(v, dict):
                    drop(v, k)

    for k in ignored_keys or ():
        drop(config, k)


def prune_configs(configs, ignored_keys=()):
    for c in configs:
        prune_config(c, ignored_keys)
    return remove_dupe_dicts(configs)


def get_tuning_keys(tuning_space):
    """Names of the parameters whose value is a list with more than one choice."""
    keys = []
    for k, v in tuning_space.items():
        if isinstance(v, dict):
            keys.extend(get_tuning_keys(v))
        if isinstance(v, list) and len(v) > 1:
            keys.append(k)
    return keys
