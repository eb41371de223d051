"""Process-group factory: DP, TP/MP, PP grid, EP + expert-DP, SP, SP x DP, hpZ intra group, a2a groups.

Reference behaviour: ``deepspeed/utils/groups.py`` (916 LoC; DP/MP :88-172, expert groups ``ep_size_N``
naming + E+D / D+E orderings :304-596, qgZ local/global all-to-all groups :672-698, SP groups
:701-852, hpZ secondary group :859-887). This is a fresh design: one rank-grid
(``pp x dp x sp x tp`` with tp fastest) created once; every other group is derived from it and cached.
On a single NVSwitch domain "node" == the whole box, so hpZ/qgZ "intra-node" groups default to the
shard group itself.
"""
import itertools

from .. import comm as dist
from .logging import log_dist

_STATE = {
    "initialized": False,
    "world": 1,
    "tp": 1,
    "pp": 1,
    "sp": 1,
    "dp": 1,
    "groups": {},  # name -> (ProcessGroup, ranks)
    "expert": {},  # "ep_size_N" -> (group, ranks)
    "expert_dp": {},
    "mpu": None,
}
expert_tensor_parallel_world_size = 1


def _ws():
    return dist.get_world_size() if dist.is_initialized() else 1


def _rk():
    return dist.get_rank() if dist.is_initialized() else 0


def _new_group_from_lists(rank_lists):
    """Create one group per rank list (all ranks must call new_group for all lists); return mine."""
    mine, mine_ranks = None, None
    me = _rk()
    for ranks in rank_lists:
        ranks = list(ranks)
        g = dist.new_group(ranks) if dist.is_initialized() and _ws() > 1 else None
        if me in ranks:
            mine, mine_ranks = g, ranks
    return mine, mine_ranks


def reset():
    _STATE.update(initialized=False, groups={}, expert={}, expert_dp={}, mpu=None, tp=1, pp=1, sp=1, dp=1)


def initialize(tp_size=1, pp_size=1, sp_size=1, mpu=None, force=False):
    """Build the rank grid. Layout (slowest -> fastest): pp, dp, sp, tp."""
    if _STATE["initialized"] and not force:
        return
    world = _ws()
    if mpu is not None:
        _STATE["mpu"] = mpu
        tp_size = mpu.get_model_parallel_world_size() if hasattr(mpu, "get_model_parallel_world_size") else tp_size
    assert world % (tp_size * pp_size * sp_size) == 0, \
        f"world {world} not divisible by tp*pp*sp = {tp_size}*{pp_size}*{sp_size}"
    dp_size = world // (tp_size * pp_size * sp_size)
    _STATE.update(initialized=True, world=world, tp=tp_size, pp=pp_size, sp=sp_size, dp=dp_size, groups={})

    def rank_of(p, d, s, t):
        return ((p * dp_size + d) * sp_size + s) * tp_size + t

    P, D, S, T = range(pp_size), range(dp_size), range(sp_size), range(tp_size)
    specs = {
        "tp": [[rank_of(p, d, s, t) for t in T] for p, d, s in itertools.product(P, D, S)],
        "sp": [[rank_of(p, d, s, t) for s in S] for p, d, t in itertools.product(P, D, T)],
        "dp": [[rank_of(p, d, s, t) for d in D] for p, s, t in itertools.product(P, S, T)],
        "pp": [[rank_of(p, d, s, t) for p in P] for d, s, t in itertools.product(D, S, T)],
        # ZeRO shards over SP x DP (reference engine.py:1839 seq_data_parallel_group)
        "sp_dp": [[rank_of(p, d, s, t) for d, s in itertools.product(D, S)] for p, t in itertools.product(P, T)],
        # model-parallel = everything that holds a different slice of the model on the same data
        "mp": [[rank_of(p, d, s, t) for p, t in itertools.product(P, T)] for d, s in itertools.product(D, S)],
    }
    for name, lists in specs.items():
        if world == 1:
            _STATE["groups"][name] = (None, [0])
            continue
        if all(len(l) == world for l in lists):
            _STATE["groups"][name] = (dist.get_world_group(), list(range(world)))
        elif all(len(l) == 1 for l in lists):
            _STATE["groups"][name] = (_single_rank_group(), [_rk()])
        else:
            _STATE["groups"][name] = _new_group_from_lists(lists)
    log_dist(f"groups: world={world} pp={pp_size} dp={dp_size} sp={sp_size} tp={tp_size}", [0])


_SINGLE = {}


def _single_rank_group():
    """A 1-rank group containing only this rank (all ranks participate in creation)."""
    if "g" not in _SINGLE:
        mine = None
        for r in range(_ws()):
            g = dist.new_group([r])
            if r == _rk():
                mine = g
        _SINGLE["g"] = mine
    return _SINGLE["g"]


def _ensure():
    if not _STATE["initialized"]:
        initialize()


def _get(name):
    _ensure()
    return _STATE["groups"][name]


# ---- reference-compatible accessors -------------------------------------------------------------
def _get_data_parallel_group():
    mpu = _STATE["mpu"]
    if mpu is not None and hasattr(mpu, "get_data_parallel_group"):
        return mpu.get_data_parallel_group()
    return _get("dp")[0]


def _get_data_parallel_world_size():
    return len(_get("dp")[1])


def _get_data_parallel_rank():
    return _get("dp")[1].index(_rk())


def _get_model_parallel_group():
    mpu = _STATE["mpu"]
    if mpu is not None and hasattr(mpu, "get_model_parallel_group"):
        return mpu.get_model_parallel_group()
    return _get("tp")[0]


def _get_model_parallel_world_size():
    return len(_get("tp")[1])


def _get_model_parallel_rank():
    return _get("tp")[1].index(_rk())


_MPU_TP_WORLD_SIZE = None  # overrides used by code that fakes a tensor-parallel context (reference groups.py:197,215)
_MPU_TP_RANK = None


def set_tensor_model_parallel_world_size(world_size):
    global _MPU_TP_WORLD_SIZE
    _MPU_TP_WORLD_SIZE = world_size


def set_tensor_model_parallel_rank(rank):
    global _MPU_TP_RANK
    _MPU_TP_RANK = rank


def get_tensor_model_parallel_world_size():
    return _MPU_TP_WORLD_SIZE if _MPU_TP_WORLD_SIZE is not None else _get_model_parallel_world_size()


def get_tensor_model_parallel_rank():
    return _MPU_TP_RANK if _MPU_TP_RANK is not None else _get_model_parallel_rank()


get_tensor_model_parallel_group = _get_model_parallel_group
get_model_parallel_group = _get_model_parallel_group
get_model_parallel_world_size = get_tensor_model_parallel_world_size
get_model_parallel_rank = get_tensor_model_parallel_rank
get_data_parallel_group = _get_data_parallel_group
get_data_parallel_world_size = _get_data_parallel_world_size
get_data_parallel_rank = _get_data_parallel_rank


def get_tensor_model_parallel_src_rank():
    return _get("tp")[1][0]


def _get_sequence_parallel_group():
    return _get("sp")[0]


def _get_sequence_parallel_world_size():
    return len(_get("sp")[1])


def _get_sequence_parallel_rank():
    return _get("sp")[1].index(_rk())


def _get_sequence_data_parallel_group():
    return _get("sp_dp")[0]


def _get_sequence_data_parallel_world_size():
    return len(_get("sp_dp")[1])


def _get_sequence_data_parallel_rank():
    return _get("sp_dp")[1].index(_rk())


def _get_pipe_parallel_group():
    return _get("pp")[0]


def _get_pipe_parallel_world_size():
    return len(_get("pp")[1])


def _get_pipe_parallel_rank():
    return _get("pp")[1].index(_rk())


def get_group_ranks(name):
    return list(_get(name)[1])


def _get_broadcast_src_rank():
    return _get("dp")[1][0]


def _get_sequence_parallel_src_rank():
    return _get("sp")[1][0]


# ---- expert parallel ---------------------------------------------------------------------------
def _expert_rank_lists(world, ep_size, mp_size=1, use_data_before_expert_parallel_=False):
    """Return (expert_parallel_lists, expert_data_parallel_lists) over the DP dimension.

    E+D (default): consecutive dp ranks form an EP group; D+E: strided.
    """
    dp = world // mp_size
    assert dp % ep_size == 0, f"dp world {dp} not divisible by ep_size {ep_size}"
    ep_lists, edp_lists = [], []
    for m in range(mp_size):
        dp_ranks = [d * mp_size + m for d in range(dp)]
        if not use_data_before_expert_parallel_:
            for i in range(dp // ep_size):
                ep_lists.append(dp_ranks[i * ep_size:(i + 1) * ep_size])
            for i in range(ep_size):
                edp_lists.append(dp_ranks[i::ep_size])
        else:
            stride = dp // ep_size
            for i in range(stride):
                ep_lists.append(dp_ranks[i::stride])
            for i in range(ep_size):
                edp_lists.append(dp_ranks[i * stride:(i + 1) * stride])
    return ep_lists, edp_lists


def _create_expert_and_data_parallel(expert_parallel_size_, use_data_before_expert_parallel_=False):
    """Create EP and expert-DP groups named ``ep_size_{N}`` (idempotent)."""
    _ensure()
    name = f"ep_size_{expert_parallel_size_}"
    if name in _STATE["expert"]:
        return
    world = _ws()
    mp = _STATE["tp"] * _STATE["pp"]
    ep_lists, edp_lists = _expert_rank_lists(world, expert_parallel_size_, mp, use_data_before_expert_parallel_)
    if world == 1:
        _STATE["expert"][name] = (None, [0])
        _STATE["expert_dp"][name] = (None, [0])
        return
    _STATE["expert"][name] = _new_group_from_lists(ep_lists) if expert_parallel_size_ > 1 else \
        (_single_rank_group(), [_rk()])
    if all(len(l) == 1 for l in edp_lists):
        _STATE["expert_dp"][name] = (_single_rank_group(), [_rk()])
    else:
        _STATE["expert_dp"][name] = _new_group_from_lists(edp_lists)
    log_dist(f"Created expert groups {name}: ep={ep_lists[:2]}... edp={edp_lists[:2]}...", [0])


_create_expert_and_data_parallel_groups = _create_expert_and_data_parallel


def _get_max_expert_size():
    assert _STATE["expert"], "Warning! Process group not initialized"
    return max(int(k.split("_")[-1]) for k in _STATE["expert"])


def _get_max_expert_size_name():
    return f"ep_size_{_get_max_expert_size()}"


def _get_max_expert_parallel_group():
    return _get_expert_parallel_group(_get_max_expert_size_name())


def _get_expert_parallel_group(group_name):
    assert group_name in _STATE["expert"], f"expert parallel group {group_name} is not initialized"
    return _STATE["expert"][group_name][0]


def _get_expert_parallel_group_dict():
    return {k: v[0] for k, v in _STATE["expert"].items()}


def _get_expert_data_parallel_group(group_name):
    assert group_name in _STATE["expert_dp"], f"expert data parallel group {group_name} is not initialized"
    return _STATE["expert_dp"][group_name][0]


def _get_expert_data_parallel_group_dict():
    return {k: v[0] for k, v in _STATE["expert_dp"].items()}


def _get_expert_parallel_world_size(group_name):
    return len(_STATE["expert"][group_name][1])


def _get_expert_parallel_rank(group_name):
    return _STATE["expert"][group_name][1].index(_rk())


def _get_expert_parallel_ranks(group_name):
    return list(_STATE["expert"][group_name][1])


def _get_expert_data_parallel_world_size(group_name):
    return len(_STATE["expert_dp"][group_name][1])


def _get_expert_data_parallel_rank(group_name):
    return _STATE["expert_dp"][group_name][1].index(_rk())


def _get_expert_broadcast_src_rank(group_name):
    return _STATE["expert_dp"][group_name][1][0]


# ---- ZeRO++ secondary (hpZ) partition and qgZ local/global all-to-all groups --------------------
def _create_zero_param_parallel_group(group_size):
    """hpZ: secondary partition group of ``group_size`` consecutive ranks inside the shard group."""
    _ensure()
    key = f"hpz_{group_size}"
    if key in _STATE["groups"]:
        return _STATE["groups"][key][0]
    base = _get("sp_dp")[1]
    world = _ws()
    assert len(base) % group_size == 0, f"shard group {len(base)} not divisible by hpz size {group_size}"
    if world == 1:
        _STATE["groups"][key] = (None, [0])
        return None
    # build for every sp_dp group in the grid
    all_lists = []
    tp, pp, sp, dp = _STATE["tp"], _STATE["pp"], _STATE["sp"], _STATE["dp"]
    for p in range(pp):
        for t in range(tp):
            ranks = [((p * dp + d) * sp + s) * tp + t for d in range(dp) for s in range(sp)]
            all_lists += [ranks[i:i + group_size] for i in range(0, len(ranks), group_size)]
    _STATE["groups"][key] = _new_group_from_lists(all_lists)
    return _STATE["groups"][key][0]


def _get_zero_param_intra_parallel_group():
    for k, v in _STATE["groups"].items():
        if k.startswith("hpz_"):
            return v[0]
    return None


def _get_zero_param_intra_parallel_group_ranks():
    for k, v in _STATE["groups"].items():
        if k.startswith("hpz_"):
            return list(v[1])
    return None


def _zero_param_parallel_is_initialized():
    return any(k.startswith("hpz_") for k in _STATE["groups"])


def _get_zero_param_intra_parallel_rank_in_mygroup():
    r = _get_zero_param_intra_parallel_group_ranks()
    return r.index(_rk()) if r else 0


def _get_zero_param_intra_parallel_group_world_size():
    r = _get_zero_param_intra_parallel_group_ranks()
    return len(r) if r else 1


def _get_local_all_to_all_group(devices_per_node=None):
    """qgZ: ``local_i`` = ranks of one node; ``global_j`` = same local index across nodes."""
    _ensure()
    world = _ws()
    import torch
    dpn = devices_per_node or max(1, min(world, torch.cuda.device_count() or world))
    key = f"a2a_{dpn}"
    if key not in _STATE["groups"]:
        if world == 1:
            _STATE["groups"][key] = ({"local_0": None, "global_0": None}, [0])
        else:
            nodes = world // dpn
            out = {}
            for i in range(nodes):
                ranks = list(range(i * dpn, (i + 1) * dpn))
                g = dist.new_group(ranks)
                if _rk() in ranks:
                    out[f"local_{i}"] = g
            if nodes > 1:
                for j in range(dpn):
                    ranks = list(range(j, world, dpn))
                    g = dist.new_group(ranks)
                    if _rk() in ranks:
                        out[f"global_{j}"] = g
            _STATE["groups"][key] = (out, list(range(world)))
    return _STATE["groups"][key][0]


def _clone_world_group():
    _ensure()
    if "world_clone" not in _STATE["groups"]:
        g = dist.new_group(list(range(_ws()))) if dist.is_initialized() and _ws() > 1 else None
        _STATE["groups"]["world_clone"] = (g, list(range(_ws())))
    return _STATE["groups"]["world_clone"][0]


def get_topology():
    _ensure()
    return {k: _STATE[k] for k in ("world", "pp", "dp", "sp", "tp")}
