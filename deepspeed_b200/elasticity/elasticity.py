"""Elastic training batch-size selection.

Given acceptable micro-batch sizes and a maximum global batch, choose a global batch size that is
divisible by as many GPU counts as possible, so a job can restart at a different scale without changing
convergence-relevant hyper-parameters. Behaviour follows reference ``elasticity/elasticity.py:27-383``
(v0.1: GPU-count search; v0.2: node-granular with model-parallel size), independently implemented.
"""
import json
import math
import os

from ..utils.logging import logger

DEEPSPEED_ELASTICITY_CONFIG = "DEEPSPEED_ELASTICITY_CONFIG"
LATEST_ELASTICITY_VERSION = 0.2
MINIMUM_DEEPSPEED_VERSION = "0.3.8"


class ElasticityError(Exception):
    pass


class ElasticityConfigError(ElasticityError):
    pass


class ElasticityIncompatibleWorldSize(ElasticityError):
    pass


# highly composite numbers: good multipliers because they have many divisors
_HCN = [1, 2, 4, 6, 12, 24, 36, 48, 60, 120, 180, 240, 360, 720, 840, 1260, 1680, 2520, 5040, 7560, 10080, 15120,
        20160, 25200, 27720, 45360, 50400, 55440, 83160, 110880, 166320, 221760, 277200, 332640, 498960, 554400,
        665280, 720720]


def get_candidate_batch_sizes(base_list, max_acceptable_batch_size):
    cands = set()
    for base in base_list:
        if base >= max_acceptable_batch_size:
            cands.add(base)
            continue
        best = base
        for h in _HCN:
            if h * base > max_acceptable_batch_size:
                break
            best = h * base
        cands.add(best)
    return sorted(cands)


def get_valid_gpus(batch_size, micro_batches, min_valid_gpus, max_valid_gpus):
    valid = set()
    for mb in micro_batches:
        if batch_size % mb != 0:
            continue
        max_gpus = batch_size // mb
        for i in range(1, int(math.isqrt(max_gpus)) + 1):
            if max_gpus % i == 0:
                for g in (i, max_gpus // i):
                    if min_valid_gpus <= g <= max_valid_gpus:
                        valid.add(g)
    return sorted(valid)


def get_best_candidates(candidate_batch_sizes, micro_batches, min_gpus, max_gpus, prefer_larger):
    max_valid, best_valid, best_bs = 0, None, int(min(micro_batches))
    for bs in candidate_batch_sizes:
        gpus = get_valid_gpus(bs, micro_batches, min_gpus, max_gpus)
        better = len(gpus) > max_valid or (len(gpus) == max_valid and
                                           ((prefer_larger and bs > best_bs) or (not prefer_larger and bs < best_bs)))
        if better:
            max_valid, best_valid, best_bs = len(gpus), gpus, bs
    return best_bs, best_valid


def _get_compatible_gpus_v01(micro_batches, max_acceptable_batch_size, min_gpus=None, max_gpus=None,
                             prefer_larger=True):
    min_gpus = min_gpus or 1
    max_gpus = max_gpus or max_acceptable_batch_size // min(micro_batches)
    if not all(mb <= max_acceptable_batch_size for mb in micro_batches):
        raise ValueError(f"All micro batches must be less than or equal to max_acceptable_batch_size: "
                         f"{max_acceptable_batch_size}")
    lcm = micro_batches[0]
    for mb in micro_batches[1:]:
        lcm = lcm * mb // math.gcd(lcm, mb)
    base_list = list(micro_batches) + [lcm]
    cands = get_candidate_batch_sizes(base_list, max_acceptable_batch_size)
    return get_best_candidates(cands, micro_batches, min_gpus, max_gpus, prefer_larger)


def _get_compatible_gpus_v02(micro_batches, max_acceptable_batch_size, current_num_gpus, min_gpus=None,
                             max_gpus=None, prefer_larger=True, num_gpus_per_node=1, model_parallel_size=1):
    if num_gpus_per_node % model_parallel_size != 0:
        raise ElasticityError(f"In Elasticity v0.2, number of GPUs per node:{num_gpus_per_node} should be divisible "
                              f"by model parallel size {model_parallel_size}")

    def get_microbatch(final_batch_size):
        candidate = None
        for mb in micro_batches:
            if (final_batch_size // current_num_gpus) % mb == 0:
                if candidate is None or (prefer_larger and mb > candidate) or (not prefer_larger and mb < candidate):
                    candidate = mb
        return candidate

    dp_size_per_node = num_gpus_per_node // model_parallel_size
    final_batch_size, valid_world_size = _get_compatible_gpus_v01(
        micro_batches, int(max_acceptable_batch_size / dp_size_per_node), int((min_gpus or 1) / num_gpus_per_node),
        int((max_gpus or max_acceptable_batch_size) / num_gpus_per_node), prefer_larger=prefer_larger)
    final_batch_size = int(final_batch_size) * dp_size_per_node
    valid_dp_world_size = [i * dp_size_per_node for i in valid_world_size]
    if current_num_gpus // model_parallel_size in valid_dp_world_size:
        return final_batch_size, valid_dp_world_size, get_microbatch(final_batch_size)
    current_dp_size = (current_num_gpus / num_gpus_per_node) * dp_size_per_node
    candidates = []
    for mb in micro_batches:
        bs = int(max_acceptable_batch_size // (mb * current_dp_size)) * mb * current_dp_size
        candidates.append(int(bs))
    final = max(candidates) if prefer_larger else min(candidates)
    return final, [int(current_dp_size)], get_microbatch(final)


def elasticity_enabled(ds_config: dict):
    return bool(ds_config.get("elasticity", {}).get("enabled", False))


def ensure_immutable_elastic_config(runtime_elastic_config_dict: dict):
    """The scheduler exported the elastic config it planned with; the runtime must see the same one."""
    if DEEPSPEED_ELASTICITY_CONFIG in os.environ:
        sched = json.loads(os.environ[DEEPSPEED_ELASTICITY_CONFIG])
        for key in ("max_train_batch_size", "micro_batch_sizes", "version"):
            a, b = sched.get(key), runtime_elastic_config_dict.get(key, sched.get(key))
            if a != b:
                raise ElasticityConfigError(f"Environment variable {DEEPSPEED_ELASTICITY_CONFIG} differs from the "
                                            f"runtime config on '{key}': {a} != {b}")
    else:
        logger.warning("Unable to find DEEPSPEED_ELASTICITY_CONFIG environment variable, cannot guarantee resource "
                       "scheduler will scale this job using compatible GPU counts.")


def compute_elastic_config(ds_config: dict, target_deepspeed_version: str = "0.0", world_size=0,
                           return_microbatch=False):
    """Return (final_batch_size, valid_gpus[, micro_batch]) for an elastic job."""
    if not isinstance(ds_config, dict):
        raise ValueError(f"Expected ds_config to be a dictionary but received a {type(ds_config)}")
    if "elasticity" not in ds_config:
        raise ElasticityConfigError("'elasticity' is missing from config json, please add it if running an elastic "
                                    "training job.")
    e = ds_config["elasticity"]
    if not e.get("enabled", False):
        raise ElasticityConfigError("Elasticity is disabled, please enable it ('enabled':true) if running an elastic "
                                    "training job.")
    from ..config.sections import ElasticityConfigModel
    cfg = ElasticityConfigModel(**e)
    mbs = list(cfg.micro_batch_sizes)
    if not mbs or not all(isinstance(m, int) and m > 0 for m in mbs):
        raise ElasticityConfigError(f"Elasticity expected micro_batch_sizes to be a list of positive ints: {mbs}")
    if cfg.min_gpus < 1 or cfg.max_gpus < 1 or cfg.max_gpus < cfg.min_gpus:
        raise ElasticityConfigError("Elasticity min/max gpus must be > 0 and min <= max")
    if cfg.model_parallel_size > 1 and float(cfg.version) != 0.2:
        raise ElasticityConfigError(f"Elasticity V{cfg.version} does not support model-parallel training.")
    if float(cfg.version) > LATEST_ELASTICITY_VERSION:
        raise ElasticityConfigError(f"Attempting to run elasticity version {cfg.version} but runtime only supports "
                                    f"up to {LATEST_ELASTICITY_VERSION}")
    if float(cfg.version) == 0.1:
        final_batch_size, valid_gpus = _get_compatible_gpus_v01(mbs, cfg.max_train_batch_size, cfg.min_gpus,
                                                                cfg.max_gpus, cfg.prefer_larger_batch)
        final_batch_size = int(final_batch_size)
        micro = None
        if world_size > 0:
            if world_size not in valid_gpus:
                raise ElasticityIncompatibleWorldSize(f"World size ({world_size}) is not valid with the current list "
                                                      f"of valid GPU counts: {valid_gpus}")
            for mb in sorted(set(mbs), reverse=True):
                if final_batch_size // world_size % mb == 0:
                    micro = mb
                    break
            assert micro is not None, "Unable to find divisible micro batch size"
    elif float(cfg.version) == 0.2:
        current = world_size if world_size > 0 else int(os.environ.get("WORLD_SIZE", 0))
        if current <= 0:
            raise ElasticityConfigError("Elasticity V 0.2 needs WORLD_SIZE to compute valid batch size.")
        final_batch_size, valid_gpus, micro = _get_compatible_gpus_v02(mbs, cfg.max_train_batch_size, current,
                                                                       cfg.min_gpus, cfg.max_gpus,
                                                                       cfg.prefer_larger_batch, cfg.num_gpus_per_node,
                                                                       cfg.model_parallel_size)
        final_batch_size = int(final_batch_size)
    else:
        raise NotImplementedError(f"Unable to find elastic logic for version: {cfg.version}")
    logger.info(f"Valid World Size (GPUs / Model Parallel Size): {valid_gpus}")
    if return_microbatch or (world_size > 0 and float(cfg.version) == 0.1 and return_microbatch):
        return final_batch_size, valid_gpus, micro
    if world_size > 0 and float(cfg.version) == 0.1:
        return final_batch_size, valid_gpus, micro
    return final_batch_size, valid_gpus


def apply_elasticity_to_config(param_dict, world_size):
    """Engine-side override of the batch triad when elasticity is enabled (reference config.py:730-790)."""
    ensure_immutable_elastic_config(param_dict["elasticity"])
    res = compute_elastic_config(param_dict, world_size=world_size, return_microbatch=True)
    final_batch_size, valid_gpus, micro = res
    ignore = param_dict["elasticity"].get("ignore_non_elastic_batch_info", False)
    if not ignore:
        for k in ("train_batch_size", "train_micro_batch_size_per_gpu", "gradient_accumulation_steps"):
            if k in param_dict:
                raise ElasticityConfigError(f"One or more batch related parameters were found in your ds_config ({k}). "
                                            "These parameters *will not be used* since elastic training is enabled; set "
                                            "'ignore_non_elastic_batch_info': true to silence this.")
    gas = final_batch_size // (micro * world_size)
    param_dict["train_batch_size"] = final_batch_size
    param_dict["train_micro_batch_size_per_gpu"] = micro
    param_dict["gradient_accumulation_steps"] = gas
    logger.info(f"[Elasticity] valid GPU counts: {valid_gpus}; batch={final_batch_size} micro={micro} gas={gas}")
