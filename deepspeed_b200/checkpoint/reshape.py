"""Topology-aware readers of a saved checkpoint directory and the PP x TP x DP contraction maps used by offline
conversion (reference ``checkpoint/reshape_utils.py``, ``reshape_meg_2d.py``, ``reshape_3d_utils.py``,
``zero_checkpoint.py``, ``deepspeed_checkpoint.py``, ``utils.py``).

A checkpoint written at (pp, tp, dp) can be read at any (pp', tp', dp') whose degrees divide the source's: every
target rank is assigned the list of source files it has to merge. File index convention: source rank
``dp * (pp * tp) + pp_index * tp + tp_index``."""
import os
import re
from collections import OrderedDict

import torch

from .constants import (ARGS_KEY, BASE_OPTIMIZER_STATE, BF16_ZERO_FILE_PREFIX, CHECKPOINT_INFO_KEY,
                        CHECKPOINT_PARALLEL_DIMS, CHECKPOINT_PP_DEGREE, CHECKPOINT_TP_DEGREE, DP_DIM,
                        FP16_ZERO_FILE_PREFIX, GROUP_PADDINGS, GROUP_STATE_KEY, ITERATION_KEY, LAYER_FILE_PREFIX,
                        MODEL_FILE_PREFIX, MODEL_FILE_SUFFIX, OPTIM_FILE_SUFFIX, OPTIMIZER_STATE_DICT, PARTITION_COUNT,
                        PP_DIM, TP_DIM, ZERO_FILE_PREFIX)


# ---- file names and discovery ---------------------------------------------------------------------
def get_model_ckpt_name_for_rank(base_folder, mp_rank_str):
    return os.path.join(base_folder, MODEL_FILE_PREFIX + mp_rank_str + MODEL_FILE_SUFFIX)


def get_zero_ckpt_name_for_rank(base_folder, dp_rank, mp_rank):
    return os.path.join(base_folder, f"{ZERO_FILE_PREFIX}{dp_rank}_{MODEL_FILE_PREFIX}{mp_rank:02d}{OPTIM_FILE_SUFFIX}")


def get_layer_ckpt_name_for_rank(base_folder, layer_id, tp_rank):
    return os.path.join(base_folder, f"{layer_id}-model_{tp_rank:02d}{MODEL_FILE_SUFFIX}")


def basic_folder_validation(dir):
    assert os.path.exists(dir), f"{dir} path does not exist"
    assert os.path.isdir(dir), f"{dir} is not a folder"


def get_files(dir):
    return [os.path.join(root, f) for root, _, files in os.walk(dir) for f in files]


def get_files_with_prefix(all_files, prefix):
    return sorted(f for f in all_files if os.path.basename(f).startswith(prefix))


def get_zero_files(dir):
    """ZeRO optimizer files sorted by (dp rank, mp rank); the three historical prefixes are tried in turn."""
    files = get_files(dir)
    for prefix in (ZERO_FILE_PREFIX, FP16_ZERO_FILE_PREFIX, BF16_ZERO_FILE_PREFIX):
        hit = get_files_with_prefix(files, prefix)
        if hit:
            rx = re.compile(f"{prefix}([0-9]+)_{MODEL_FILE_PREFIX}([0-9]+)")
            keyed = []
            for f in hit:
                m = rx.search(f)
                if not m:
                    raise ValueError(f"Cannot parse dp_rank and mp_rank from {f}")
                keyed.append((int(m.group(1)), int(m.group(2)), f))
            return [f for _, _, f in sorted(keyed)]
    return []


def partition_data(data_list, num_partitions):
    n = len(data_list)
    assert n % num_partitions == 0, f"{n} items cannot be split into {num_partitions} equal partitions"
    size = n // num_partitions
    return [data_list[i:i + size] for i in range(0, n, size)]


def merge_state(state_a, state_b, key_list=()):
    """Merge two rank states of the same structure: tensors are concatenated on dim 0 (flat partitions), containers
    merged element-wise, everything else taken from ``state_a``."""
    if type(state_a) is not type(state_b):
        raise ValueError(f"Cannot merge two states of types {type(state_a)} and type {type(state_b)} "
                         f"(at {'.'.join(map(str, key_list))})")
    if isinstance(state_a, (dict, OrderedDict)):
        out = type(state_a)()
        for k, v in state_b.items():
            out[k] = merge_state(state_a[k], v, list(key_list) + [k]) if k in state_a else v
        return out
    if isinstance(state_a, (list, tuple)):
        if len(state_a) != len(state_b):
            raise ValueError(f"Cannot merge lists of different lengths, a = {len(state_a)} b = {len(state_b)} "
                             f"(at {'.'.join(map(str, key_list))})")
        return type(state_a)(merge_state(a, b, key_list) for a, b in zip(state_a, state_b))
    if torch.is_tensor(state_a):
        return torch.cat([state_a, state_b], 0)
    return state_a


# ---- 2-D (pp x tp) and 3-D (+ dp) maps -------------------------------------------------------------
class meg_2d_parallel_map:
    """(pp index, tp index) -> list of source items."""

    def __init__(self, pp_degree, tp_degree):
        self.pp_degree, self.tp_degree = pp_degree, tp_degree
        self.map = {}

    def simple_init(self):
        self.map = {f"{i},{j}": [i * self.tp_degree + j] for i in range(self.pp_degree)
                    for j in range(self.tp_degree)}

    def add_data(self, pp_index, tp_index, data):
        assert pp_index < self.pp_degree and tp_index < self.tp_degree
        assert isinstance(data, list)
        self.map.setdefault(f"{pp_index},{tp_index}", []).extend(data)

    def get_data(self, pp_index=None, tp_index=None):
        pps = range(self.pp_degree) if pp_index is None else [pp_index]
        tps = range(self.tp_degree) if tp_index is None else [tp_index]
        out = []
        for i in pps:
            for j in tps:
                out += self.map[f"{i},{j}"]
        return out

    def print_data(self, tag):
        print(tag)
        for k, v in self.map.items():
            print(f"{k} = {v}")


def reshape_meg_2d_parallel(old_pp_degree, old_tp_degree, new_pp_degree, new_tp_degree, verbose=False):
    assert new_pp_degree <= old_pp_degree and new_tp_degree <= old_tp_degree
    cur = meg_2d_parallel_map(old_pp_degree, old_tp_degree)
    cur.simple_init()
    if new_tp_degree != old_tp_degree:
        nxt = meg_2d_parallel_map(old_pp_degree, new_tp_degree)
        for i in range(old_pp_degree):
            for j, part in enumerate(partition_data(cur.get_data(pp_index=i), new_tp_degree)):
                nxt.add_data(i, j, part)
        cur = nxt
    if new_pp_degree != old_pp_degree:
        nxt = meg_2d_parallel_map(new_pp_degree, cur.tp_degree)
        for j in range(cur.tp_degree):
            for i, part in enumerate(partition_data(cur.get_data(tp_index=j), new_pp_degree)):
                nxt.add_data(i, j, part)
        cur = nxt
    if verbose:
        cur.print_data("final_2d_map:")
    return cur


class model_3d_desc:

    def __init__(self, pp_degree=1, tp_degree=1, dp_degree=1):
        self.pp_degree, self.tp_degree, self.dp_degree = pp_degree, tp_degree, dp_degree

    def get_desc(self):
        return f"{PP_DIM},{TP_DIM},{DP_DIM} = ({self.pp_degree}, {self.tp_degree}, {self.dp_degree})"

    def world_size(self):
        return self.pp_degree * self.tp_degree * self.dp_degree

    def is_valid(self, pp_index, tp_index, dp_index):
        errs = [f"{name} indexing error: index {i} >= degree {d}"
                for i, d, name in ((pp_index, self.pp_degree, PP_DIM), (tp_index, self.tp_degree, TP_DIM),
                                   (dp_index, self.dp_degree, DP_DIM)) if i >= d]
        return not errs, errs

    def can_reshape(self, target):
        errs = [f"Expansion reshape not supported - {name}: {a} ---> {b}"
                for a, b, name in ((self.pp_degree, target.pp_degree, PP_DIM), (self.tp_degree, target.tp_degree, TP_DIM),
                                   (self.dp_degree, target.dp_degree, DP_DIM)) if b > a]
        return not errs, errs

    def reshape(self, target, verbose=False):
        """List (one 2-D map per target dp index) of the source file indices every target rank merges."""
        ok, errs = self.can_reshape(target)
        assert ok, ",".join(errs)
        m2d = reshape_meg_2d_parallel(self.pp_degree, self.tp_degree, target.pp_degree, target.tp_degree, verbose)
        per_dp = [meg_2d_parallel_map(m2d.pp_degree, m2d.tp_degree) for _ in range(target.dp_degree)]
        src_2d = self.pp_degree * self.tp_degree
        for i in range(m2d.pp_degree):
            for j in range(m2d.tp_degree):
                flat = []
                for idx in m2d.get_data(i, j):
                    flat += [idx + d * src_2d for d in range(self.dp_degree)]
                for part, tgt in zip(partition_data(flat, target.dp_degree), per_dp):
                    tgt.add_data(i, j, part)
        return per_dp


def _descriptor_from_metadata(file_list, zero_file_list, model_state_metadata):
    model_files = [f for f, _ in model_state_metadata]
    dims = [(f, md.get(CHECKPOINT_PARALLEL_DIMS)) for f, md in model_state_metadata]
    have = [(f, d) for f, d in dims if d is not None]
    if have:
        if len(have) != len(dims):
            raise RuntimeError(f"Checkpoint {CHECKPOINT_PARALLEL_DIMS} is missing from model-state files: "
                               f"{[f for f, d in dims if d is None]}")
        vals = set()
        for f, d in have:
            if not isinstance(d, dict):
                raise RuntimeError(f"Checkpoint {CHECKPOINT_PARALLEL_DIMS} must be a dictionary in {f}.")
            pair = (d.get(CHECKPOINT_PP_DEGREE), d.get(CHECKPOINT_TP_DEGREE))
            for key, v in zip((CHECKPOINT_PP_DEGREE, CHECKPOINT_TP_DEGREE), pair):
                if isinstance(v, bool) or not isinstance(v, int) or v < 1:
                    raise RuntimeError(f"Checkpoint {key} must be a positive integer in {f}, got {v!r}.")
            vals.add(pair)
        if len(vals) != 1:
            raise RuntimeError(f"Checkpoint {CHECKPOINT_PARALLEL_DIMS} disagrees across model-state files: {vals}")
        pp, tp = vals.pop()
        if len(model_files) != pp * tp:
            raise RuntimeError(f"Checkpoint model-state file count {len(model_files)} does not match "
                               f"{CHECKPOINT_PP_DEGREE}={pp} * {CHECKPOINT_TP_DEGREE}={tp}.")
        if len(zero_file_list) % (pp * tp) != 0:
            raise RuntimeError(f"Checkpoint ZeRO file count {len(zero_file_list)} is not divisible by {pp * tp}.")
        return model_3d_desc(pp, tp, max(1, len(zero_file_list) // (pp * tp)))
    # older checkpoints: infer from the file population (layer_01-* files exist only with pipeline parallelism)
    n_pp0 = len(get_files_with_prefix(file_list, f"{LAYER_FILE_PREFIX}01"))
    if n_pp0 > 0:
        tp = n_pp0
        pp = len(model_files) // tp
    else:
        tp, pp = len(model_files), 1
    return model_3d_desc(pp, tp, max(1, len(zero_file_list) // (pp * tp)))


def get_model_3d_descriptor(dir):
    files = get_files(dir)
    meta = []
    for f in get_files_with_prefix(files, MODEL_FILE_PREFIX):
        sd = torch.load(f, map_location="cpu", weights_only=False)
        meta.append((f, {CHECKPOINT_PARALLEL_DIMS: sd.get(CHECKPOINT_PARALLEL_DIMS)}))
        del sd
    return _descriptor_from_metadata(files, get_zero_files(dir), meta)


# ---- readers --------------------------------------------------------------------------------------
class ZeROCheckpoint:
    """The ZeRO optimizer files of a checkpoint directory addressed by target (pp, tp, dp) coordinates."""

    def __init__(self, dir, source_3d=None):
        basic_folder_validation(dir)
        self.dir = dir
        self.file_list = get_zero_files(dir)
        self.num_files = len(self.file_list)
        assert self.num_files > 0, f"No ZeRO files found in {dir}"
        self.src_3d = get_model_3d_descriptor(dir) if source_3d is None else source_3d
        self.target_3d = model_3d_desc(self.src_3d.pp_degree, self.src_3d.tp_degree, self.src_3d.dp_degree)
        self._3d_file_map = self.src_3d.reshape(self.target_3d)

    def get_src_world_size(self):
        return self.src_3d.world_size()

    def get_src_tp_degree(self):
        return self.src_3d.tp_degree

    def get_src_pp_degree(self):
        return self.src_3d.pp_degree

    def get_src_dp_degree(self):
        return self.src_3d.dp_degree

    def reshape(self, target_3d_desc):
        self.target_3d = target_3d_desc
        self._3d_file_map = self.src_3d.reshape(target_3d_desc)

    def get_file_indices_for_rank(self, pp_index, tp_index, dp_index):
        assert dp_index < len(self._3d_file_map), f"DP index {dp_index} >= DP degree {len(self._3d_file_map)}"
        return self._3d_file_map[dp_index].get_data(pp_index, tp_index)

    def get_files_for_rank(self, pp_index, tp_index, dp_index):
        return [self.file_list[i] for i in self.get_file_indices_for_rank(pp_index, tp_index, dp_index)]

    def get_state_for_rank(self, pp_index, tp_index, dp_index, keys_to_ignore=(), strip_tensor_paddings=True):
        merged = None
        for path in self.get_files_for_rank(pp_index, tp_index, dp_index):
            sd = torch.load(path, map_location="cpu", weights_only=False)
            for k in keys_to_ignore:
                sd.pop(k, None)
            if strip_tensor_paddings:
                self._strip_paddings(sd)
            merged = sd if merged is None else merge_state(merged, sd)
            opt = merged.get(OPTIMIZER_STATE_DICT)
            if opt and opt.get(PARTITION_COUNT):
                opt[PARTITION_COUNT] = [self.target_3d.dp_degree] * len(opt[PARTITION_COUNT])
            if strip_tensor_paddings and opt and opt.get(GROUP_PADDINGS):
                opt[GROUP_PADDINGS] = [0] * len(opt[GROUP_PADDINGS])
        return merged

    @staticmethod
    def _strip_paddings(sd):
        opt = sd.get(OPTIMIZER_STATE_DICT) or {}
        pads = opt.get(GROUP_PADDINGS)
        states = (opt.get(BASE_OPTIMIZER_STATE) or {}).get(GROUP_STATE_KEY) if isinstance(
            opt.get(BASE_OPTIMIZER_STATE), dict) else None
        if not pads or not states:
            return
        for key, group_state in states.items():
            if pads[key] == 0:
                continue
            for name, val in group_state.items():
                if name != "step" and torch.is_tensor(val) and val.numel() > pads[key]:
                    group_state[name] = val.narrow(0, 0, val.numel() - pads[key]).clone()

    def print_3d_index_map(self, tag=None):
        if tag:
            print(f"3D index map: {tag}")
        for d, m in enumerate(self._3d_file_map):
            m.print_data(f"dp = {d}")

    def print_3d_file_map(self, tag=None):
        if tag:
            print(f"3D file map: {tag}")
        for d, m in enumerate(self._3d_file_map):
            for i in range(m.pp_degree):
                for j in range(m.tp_degree):
                    print(f"{i}, {j}, {d} => {[self.file_list[x] for x in m.get_data(i, j)]}")


class DeepSpeedCheckpoint:
    """A saved (Megatron-style PP x TP x DP) checkpoint directory seen at a possibly smaller target topology: which
    model-state / layer / ZeRO files each target rank reads, plus their merged contents."""

    def __init__(self, dir, tp_degree=None, pp_degree=None, dp_degree=None, final_layer_norm_idx=-2):
        basic_folder_validation(dir)
        self.dir = dir
        self.file_list = get_files(dir)
        self.mp_rank_files = get_files_with_prefix(self.file_list, MODEL_FILE_PREFIX)
        self.layer_files = sorted(f for f in get_files_with_prefix(self.file_list, LAYER_FILE_PREFIX)
                                  if "-model_" in os.path.basename(f))
        self.zero_files = get_zero_files(dir)
        self.zero_checkpoint = ZeROCheckpoint(dir) if self.zero_files else None
        self.src_3d = self.zero_checkpoint.src_3d if self.zero_checkpoint else get_model_3d_descriptor(dir)
        self.original_tp_degree, self.original_pp_degree = self.src_3d.tp_degree, self.src_3d.pp_degree
        self.original_dp_degree = self.src_3d.dp_degree
        self.tp_degree = self.original_tp_degree if tp_degree is None else tp_degree
        self.pp_degree = self.original_pp_degree if pp_degree is None else pp_degree
        self.dp_degree = self.original_dp_degree if dp_degree is None else dp_degree
        self.original_world_size = self.src_3d.world_size()
        self.world_size = self.tp_degree * self.pp_degree * self.dp_degree
        self.final_layer_norm_idx = final_layer_norm_idx
        self.old_2d_map = meg_2d_parallel_map(self.original_pp_degree, self.original_tp_degree)
        self.old_2d_map.simple_init()
        self.new_2d_map = reshape_meg_2d_parallel(self.original_pp_degree, self.original_tp_degree, self.pp_degree,
                                                  self.tp_degree)
        if self.zero_checkpoint is not None and self.is_change_dp_degree() or self.is_change_pp_degree() \
                or self.is_change_tp_degree():
            if self.zero_checkpoint is not None:
                self.zero_checkpoint.reshape(model_3d_desc(self.pp_degree, self.tp_degree, self.dp_degree))
        self.global_state = {}
        self.layer_keys = self._layer_keys()
        self.layer_count = len(self.layer_keys)

    def is_change_tp_degree(self):
        return self.tp_degree != self.original_tp_degree

    def is_change_pp_degree(self):
        return self.pp_degree != self.original_pp_degree

    def is_change_dp_degree(self):
        return self.dp_degree != self.original_dp_degree

    def _layer_keys(self):
        keys = set()
        for f in self.layer_files:
            m = re.match(rf"({LAYER_FILE_PREFIX}\\d+)-model_", os.path.basename(f))
            if m:
                keys.add(m.group(1))
        return sorted(keys)

    def validate_files(self):
        for f in self.file_list:
            if not os.path.isfile(f):
                print(f"Error: {f} is not existent")

    def show_2d_mapping(self):
        self.old_2d_map.print_data("reshaped 2d map ---- begin")
        self.new_2d_map.print_data("reshaped 2d map ---- end")

    # ---- ZeRO state ----
    def get_zero_checkpoint_state(self, pp_index, tp_index, dp_index, strip_tensor_paddings=True):
        return self.zero_checkpoint.get_state_for_rank(pp_index, tp_index, dp_index, keys_to_ignore=[PARTITION_COUNT],
                                                       strip_tensor_paddings=strip_tensor_paddings)

    def get_zero_files(self, pp_index, tp_index, dp_index):
        return self.zero_checkpoint.get_files_for_rank(pp_index, tp_index, dp_index)

    # ---- model-state files ----
    def get_2d_parallel_files(self, tp_index, pp_index):
        return [self.mp_rank_files[i] for i in self.new_2d_map.get_data(pp_index=pp_index, tp_index=tp_index)]

    def get_2d_parallel_state(self, tp_index, pp_index):
        merged = None
        for f in self.get_2d_parallel_files(tp_index, pp_index):
            sd = torch.load(f, map_location="cpu", weights_only=False)
            merged = sd if merged is None else merge_state(merged, sd)
        return merged

    def _checkpoint_value(self, key):
        if key not in self.global_state and self.mp_rank_files:
            sd = torch.load(self.mp_rank_files[0], map_location="cpu", weights_only=False)
            self.global_state[key] = sd.get(key)
        return self.global_state.get(key)

    def get_iteration(self):
        return self._checkpoint_value(ITERATION_KEY) or 0

    def get_args(self):
        return self._checkpoint_value(ARGS_KEY)

    def get_checkpoint_info(self, info_key=CHECKPOINT_INFO_KEY):
        return self._checkpoint_value(info_key)

    # ---- per-layer files (pipeline checkpoints) ----
    def _files_of_layer(self, layer_key, tp_index=None):
        files = [f for f in self.layer_files if os.path.basename(f).startswith(layer_key + "-")]
        if tp_index is None:
            return files
        per_tp = partition_data(files, self.tp_degree) if files else []
        return per_tp[tp_index] if per_tp else []

    def _merged_layer(self, layer_key, tp_index):
        sds = [torch.load(f, map_location="cpu", weights_only=False) for f in self._files_of_layer(layer_key, tp_index)]
        if not sds:
            return {}
        out = OrderedDict()
        for k in sds[0]:
            vals = [sd[k] for sd in sds]
            out[k] = vals[0] if len(vals) == 1 or not torch.is_tensor(vals[0]) else vals
        return out

    def get_embedding_layer_id(self):
        return self.layer_keys[0]

    def get_final_norm_layer_id(self):
        return self.layer_keys[self.final_layer_norm_idx]

    def get_embedding_files(self, tp_index):
        return self._files_of_layer(self.get_embedding_layer_id(), tp_index)

    def get_embedding_state(self, tp_index):
        return self._merged_layer(self.get_embedding_layer_id(), tp_index)

    def get_final_norm_files(self, tp_index):
        return self._files_of_layer(self.get_final_norm_layer_id(), tp_index)

    def get_final_norm_state(self, tp_index):
        files = self.get_final_norm_files(tp_index)
        return torch.load(files[0], map_location="cpu", weights_only=False) if files else {}

    def get_pp_transformer_map(self, pp_index):
        body = self.layer_keys[1:self.final_layer_norm_idx]
        return partition_data(body, self.pp_degree)[pp_index] if body else []

    def get_transformer_state(self, tp_index, pp_index):
        return [self._merged_layer(k, tp_index) for k in self.get_pp_transformer_map(pp_index)]
