"""Engine checkpoint save/load in the reference on-disk layout.

Files and dictionary keys follow ``deepspeed/runtime/engine.py`` (``_get_ckpt_name`` :4212-4236,
``_get_zero_ckpt_name`` :4195-4205, ``_common_checkpoint_state`` :4897-4919, ``save_checkpoint`` :4749-4846,
``load_checkpoint`` :4271-4366, tag validation :4732-4747) so that the reference's ``zero_to_fp32.py`` and
``ds_to_universal.py`` read our checkpoints and vice versa. The shard layout translation lives in
:mod:`.layout`.
"""
import hashlib
import os
import shutil
from collections import OrderedDict

import torch

from .. import comm as dist
from ..utils.logging import log_dist, logger
from ..version import __version__
from .layout import ReferenceLayout

BUFFER_NAMES = "buffer_names"
PARAM_SHAPES = "param_shapes"
FROZEN_PARAM_SHAPES = "frozen_param_shapes"
FROZEN_PARAM_FRAGMENTS = "frozen_param_fragments"
OPTIMIZER_STATE_DICT = "optimizer_state_dict"
FP32_FLAT_GROUPS = "fp32_flat_groups"
SINGLE_PARTITION_OF_FP32_GROUPS = "single_partition_of_fp32_groups"
BASE_OPTIMIZER_STATE = "base_optimizer_state"
UNIVERSAL_CHECKPOINT_INFO = "universal_checkpoint_info"


# ---------------------------------------------------------------------------------------------------------
# names
# ---------------------------------------------------------------------------------------------------------
def _mp_rank(engine):
    from ..utils import groups
    return groups._get_model_parallel_rank() if engine.mp_world_size > 1 else 0


def get_ckpt_name(engine, ckpt_dir, tag, mp_placeholder=None):
    mp = f"{_mp_rank(engine):02d}" if mp_placeholder is None else mp_placeholder
    if engine.zero_optimization_partition_weights():
        fname = f"zero_pp_rank_{engine.seq_dp_rank}_mp_rank_{mp}_model_states.pt"
    else:
        fname = f"mp_rank_{mp}_model_states.pt"
    return os.path.join(ckpt_dir, str(tag), fname)


def get_zero_ckpt_name(engine, ckpt_dir, tag, dp_rank=None):
    dp_rank = engine.seq_dp_rank if dp_rank is None else dp_rank
    prefix = "bf16_" if engine.bfloat16_enabled() else ""
    return os.path.join(ckpt_dir, str(tag), f"{prefix}zero_pp_rank_{dp_rank}_mp_rank_{_mp_rank(engine):02d}_optim_states.pt")


def get_expert_ckpt_name(ckpt_dir, layer_id, expert_id, tag, mp_rank=0):
    return os.path.join(ckpt_dir, str(tag), f"layer_{layer_id}_expert_{expert_id}_mp_rank_{mp_rank:02d}_model_states.pt")


# ---------------------------------------------------------------------------------------------------------
# tag validation (all ranks must agree on the tag)
# ---------------------------------------------------------------------------------------------------------
def _checkpoint_tag_validation(engine, tag):
    if not engine.checkpoint_tag_validation_enabled() or dist.get_world_size() == 1:
        return
    s_hash = hashlib.sha1(str(tag).encode())
    bhash = torch.tensor(list(s_hash.digest()), dtype=torch.int64, device=engine.device)
    mx, mn = bhash.clone(), bhash.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    valid = bool(torch.equal(bhash, mx) and torch.equal(bhash, mn))
    msg = (f"[rank={dist.get_rank()}] The checkpoint tag name '{tag}' is not consistent across all ranks. "
           "Including rank unique information in checkpoint tag could cause issues when restoring with different "
           "world sizes.")
    if engine.checkpoint_tag_validation_fail():
        assert valid, msg
    elif not valid:
        logger.warning(msg)


# ---------------------------------------------------------------------------------------------------------
# save
# ---------------------------------------------------------------------------------------------------------
def _shared_params(module):
    """name -> canonical name for parameters registered under several names (tied weights)."""
    first, shared = {}, {}
    for name, p in module.named_parameters(remove_duplicate=False):
        if id(p) in first:
            shared[name] = first[id(p)]
        else:
            first[id(p)] = name
    return shared


def _module_state_for_save(engine, exclude_frozen):
    if engine.zero_optimization_partition_weights():
        # placeholders for partitioned params (shape metadata lives in param_shapes) + real buffers
        sd = OrderedDict()
        for name, p in engine.module.named_parameters():
            u = getattr(p, "ds_unit", None)
            if u is not None and u.persistent:
                sd[name] = p.detach().clone()
            else:
                sd[name] = torch.empty(0, dtype=p.dtype)
        for name, b in engine.module.named_buffers():
            sd[name] = b
        return sd
    return engine.module_state_dict(exclude_frozen_parameters=exclude_frozen)


def _cdiv(a, b):
    return (a + b - 1) // b


@torch.no_grad()
def _zero3_frozen_fragments(engine):
    """ZeRO-3: this rank's fragment of every frozen parameter in the reference layout (``param.ds_tensor``:
    ``ceil(numel/world)`` contiguous elements of the flat parameter, zero padded on the last ranks; reference
    ``engine.py:5292-5306``), so that the reference ``zero_to_fp32._zero3_merge_frozen_params`` (cat + narrow)
    rebuilds them. Collective: one all-gather per unit that holds frozen parameters."""
    z = engine.optimizer
    shapes, frags = OrderedDict(), OrderedDict()
    by_id = {}
    for u in z.units:
        frozen = [i for i, p in enumerate(u.params) if not p.requires_grad]
        if not frozen:
            continue
        if u.world > 1 and not u.persistent and u.param_shard is not None:
            full = z._gather_shard_tensor(u, u.param_shard)
        elif u.param_shard is not None and not u.persistent:
            full = u.param_shard
        else:
            full = None
        world, rank = max(u.world, 1), (z._rank_of(u.group_key) if u.world > 1 else 0)
        for i in frozen:
            p, n, o = u.params[i], u.numels[i], u.offsets[i]
            flat = full[o:o + n] if full is not None else p.detach().reshape(-1)
            per = _cdiv(n, world)
            a, b = min(rank * per, n), min((rank + 1) * per, n)
            frag = torch.zeros(per, dtype=flat.dtype)
            if b > a:
                frag[:b - a].copy_(flat[a:b])
            by_id[id(p)] = (torch.Size(u.shapes[i]), frag)
    for name, p in engine.module.named_parameters():
        if not p.requires_grad:
            if id(p) in by_id:
                shapes[name], frags[name] = by_id[id(p)]
            else:  # not managed by a unit (should not happen at stage 3): keep the tensor itself
                shapes[name], frags[name] = torch.Size(p.shape), p.detach().cpu().clone()
    return shapes, frags


@torch.no_grad()
def _zero3_load_persistent_from_module(engine, ckpt):
    """No optimizer shard files: replicated (persistent) parameters are in ``ckpt['module']``; copy them into the
    live tensors before the flat shards are refreshed from the parameters."""
    z = engine.optimizer
    module_sd = ckpt.get("module") or {}
    names = {id(p): n for n, p in engine.module.named_parameters()}
    for u in z.units:
        if not u.persistent:
            continue
        for i, p in enumerate(u.params):
            src = module_sd.get(names.get(id(p)))
            if src is not None and src.numel() == u.numels[i] and p.numel() == u.numels[i]:
                p.data.copy_(src.to(p.device).view_as(p))


@torch.no_grad()
def _zero3_load_frozen(engine, ckpt):
    """Inverse of :func:`_zero3_frozen_fragments` (collective: one all-reduce per unit with frozen parameters).
    Writes the fp32 master region, the low-precision shard and, for resident units, the gathered copy."""
    z = engine.optimizer
    frags = ckpt.get(FROZEN_PARAM_FRAGMENTS) or {}
    if not frags:
        return
    names = {id(p): n for n, p in engine.module.named_parameters()}
    for u in z.units:
        frozen = [i for i, p in enumerate(u.params) if not p.requires_grad and names.get(id(p)) in frags]
        if not frozen:
            continue
        world, rank = max(u.world, 1), (z._rank_of(u.group_key) if u.world > 1 else 0)
        full = torch.zeros(u.numel_padded, dtype=torch.float32, device=u.device)
        for i in frozen:
            n, o = u.numels[i], u.offsets[i]
            frag = frags[names[id(u.params[i])]].reshape(-1)
            per = _cdiv(n, world)
            a, b = min(rank * per, n), min((rank + 1) * per, n)
            if frag.numel() == n and world > 1:  # unpartitioned fragment (saved by a persistent/replicated param)
                if rank == 0:
                    full[o:o + n].copy_(frag.to(u.device, torch.float32))
            elif b > a and frag.numel() >= b - a:
                full[o + a:o + b].copy_(frag[:b - a].to(u.device, torch.float32))
        if u.world > 1:
            dist.all_reduce(full, group=u.dp_group)
        lo, hi = u.shard_range(rank if u.world > 1 else 0)
        st = z.storages[(u.group_key, u.dtype)]
        for i in frozen:
            n, o = u.numels[i], u.offsets[i]
            a, b = max(o, lo), min(o + n, hi)
            if b > a:
                if u.master_shard is not None:
                    u.master_shard[a - lo:b - lo].copy_(full[a:b].to(u.master_shard.dtype))
                if u.param_shard is not None:
                    u.param_shard[a - lo:b - lo].copy_(full[a:b].to(u.param_shard.dtype))
                if st.lp is not None:
                    st.lp[u.shard_offset + a - lo:u.shard_offset + b - lo].copy_(full[a:b].to(st.lp.dtype))
            if u.full_param is not None:
                u.full_param[o:o + n].copy_(full[o:o + n].to(u.full_param.dtype))


def _zero_optimizer_state(engine):
    """Reference-layout optimizer state dict for this rank."""
    z = engine.optimizer
    lay = ReferenceLayout(z)
    fp32 = lay.export(lambda u: u.master_shard)
    state_keys = set()
    for stt in z.optimizer.state.values():
        for k, v in stt.items():
            if torch.is_tensor(v) and v.numel() > 1:
                state_keys.add(k)
    if getattr(z, "_nvme", None) is not None:  # moments swapped out to NVMe are not in optimizer.state
        state_keys.update(z._nvme_names)
    if z._shard_world > 1 and dist.is_initialized():
        # a rank whose shards hold only frozen runs has no state tensors: agree on the key set before the
        # per-key collectives below
        gathered = [None] * z._shard_world
        dist.all_gather_object(gathered, sorted(state_keys), group=z._shard_group)
        state_keys = set().union(*[set(g or ()) for g in gathered])
    exported = {k: lay.export(lambda u, k=k: z._state_shard(u, k)) for k in sorted(state_keys)}
    # per-group scalar state (step)
    base_state = {}
    n_groups = len(z.optimizer.param_groups)
    for gi in range(n_groups):
        entry = {}
        for (skey, a, b, g, u, v) in z._segments:
            if g == gi:
                for k, val in z.optimizer.state.get(v, {}).items():
                    if not (torch.is_tensor(val) and val.numel() > 1):
                        entry[k] = val
                break
        for k in exported:
            entry[k] = exported[k][gi]
        base_state[gi] = entry
    pgs = []
    for gi, g in enumerate(z.optimizer.param_groups):
        d = {k: v for k, v in g.items() if k != "params"}
        d["params"] = [gi]
        pgs.append(d)
    base = {"state": base_state, "param_groups": pgs}
    sd = {
        "loss_scaler": z.loss_scaler,
        "dynamic_loss_scale": z.dynamic_loss_scale,
        "overflow": z.overflow,
        "partition_count": z._shard_world,
        "zero_stage": z.stage,
        "ds_version": __version__,
    }
    # expert units hold different parameters on every expert-parallel rank and are sharded over their own group: they
    # do not fit the reference's data-parallel flats and are stored per unit (this rank's shard of master + moments)
    experts = OrderedDict()
    for u in z.units:
        if getattr(u, "group_key", "dp") == "dp":
            continue
        from ..utils import groups as _groups
        try:
            ep_rank, ep_size = _groups._get_expert_parallel_rank(u.group_key), \
                _groups._get_expert_parallel_world_size(u.group_key)
        except Exception:  # noqa: BLE001
            ep_rank, ep_size = 0, 1
        entry = {"names": list(u.names), "world": u.world, "rank": u.rank, "shapes": [tuple(sh) for sh in u.shapes],
                 "offsets": list(u.offsets), "numel_padded": u.numel_padded, "ep_rank": ep_rank, "ep_size": ep_size,
                 "fp32": u.master_shard.detach().float().cpu().clone(), "state": {}}
        for k in sorted(state_keys):
            st = z._state_shard(u, k)
            if st is not None:
                entry["state"][k] = st.detach().float().cpu().clone()
        experts[u.name] = entry
    if experts:
        sd["expert_units"] = experts
    if z.stage >= 3:
        sd[OPTIMIZER_STATE_DICT] = base
        sd[FP32_FLAT_GROUPS] = fp32
    else:
        sd["clip_grad"] = z.clip_grad
        sd[BASE_OPTIMIZER_STATE] = base
        # padding of the last rank is stripped on disk (reference _get_groups_without_padding)
        pads = lay.group_paddings()
        sd[SINGLE_PARTITION_OF_FP32_GROUPS] = [t[:t.numel() - pad] if pad else t for t, pad in zip(fp32, pads)]
        sd["group_paddings"] = pads
        sd["param_slice_mappings"] = lay.param_slice_mappings()
    return sd, lay


def _data_sampler_state(engine):
    sampler = engine.get_data_sampler() if hasattr(engine, "get_data_sampler") else None
    if sampler is not None and hasattr(sampler, "state_dict") and engine.curriculum_learning_enabled():
        return sampler.state_dict()
    return None


def _random_ltd_state(engine):
    sch = getattr(engine, "random_ltd_scheduler", None)
    return sch.state_dict() if sch is not None else None


def save_engine_checkpoint(engine, save_dir, tag=None, client_state=None, save_latest=True,
                           exclude_frozen_parameters=False):
    if tag is None:
        tag = f"global_step{engine.global_steps}"
    tag = str(tag)
    if hasattr(engine.optimizer, "_join_step"):
        engine.optimizer._join_step()  # a pipelined optimizer step may still be running on its side stream
    ce = engine.checkpoint_engine
    from ..runtime.checkpoint_engine import CheckpointCommitInfo
    info = CheckpointCommitInfo(tag=tag, save_dir=save_dir, save_latest=save_latest)
    ce.create(info)
    if engine.global_rank == 0:
        ce.makedirs(os.path.join(save_dir, tag), exist_ok=True)
    dist.barrier()
    _checkpoint_tag_validation(engine, tag)
    os.makedirs(os.path.join(save_dir, tag), exist_ok=True)

    lay = None
    zero_sd = None
    if engine.optimizer is not None and engine.save_zero_checkpoint:
        zero_sd, lay = _zero_optimizer_state(engine)  # collective: every rank participates

    if engine.has_moe_layers:
        from ..moe.checkpoint import save_expert_optimizer_state, save_moe_experts
        save_moe_experts(engine, save_dir, tag)
        if not engine.save_zero_checkpoint:
            save_expert_optimizer_state(engine, save_dir, tag)

    if engine.save_non_zero_checkpoint:
        module_sd = _module_state_for_save(engine, exclude_frozen_parameters)
        if engine.has_moe_layers:
            from ..moe.checkpoint import strip_expert_params
            module_sd = strip_expert_params(engine, module_sd)
        buffer_names = [n for n, _ in engine.module.named_buffers()]
        if lay is None and engine.optimizer is not None:
            lay = ReferenceLayout(engine.optimizer)
        param_shapes = lay.param_shapes() if lay is not None else None
        frozen_shapes, frozen_frags = OrderedDict(), OrderedDict()
        if exclude_frozen_parameters:
            pass
        elif engine.zero_optimization_partition_weights() and engine.optimizer is not None \
                and hasattr(engine.optimizer, "units"):
            frozen_shapes, frozen_frags = _zero3_frozen_fragments(engine)  # collective
        else:
            for n, p in engine.module.named_parameters():
                if not p.requires_grad:
                    frozen_shapes[n] = torch.Size(getattr(p, "ds_shape", p.shape))
                    frozen_frags[n] = p.detach().cpu().clone()
        state = dict(module=module_sd, buffer_names=buffer_names, optimizer=None, param_shapes=param_shapes,
                     frozen_param_shapes=frozen_shapes or None, frozen_param_fragments=frozen_frags or None,
                     shared_params=_shared_params(engine.module),
                     lr_scheduler=engine.lr_scheduler.state_dict() if engine.lr_scheduler is not None else None,
                     data_sampler=_data_sampler_state(engine), random_ltd=_random_ltd_state(engine),
                     sparse_tensor_module_names=set(getattr(engine, "sparse_tensor_module_names", ())),
                     skipped_steps=engine.skipped_steps, global_steps=engine.global_steps,
                     global_samples=engine.global_samples, dp_world_size=engine.seq_dp_world_size,
                     mp_world_size=engine.mp_world_size,
                     checkpoint_parallel_dimensions={"pp_degree": 1, "tp_degree": engine.mp_world_size},
                     ds_config=engine.config, ds_version=__version__)
        if engine.optimizer is not None and not engine.save_zero_checkpoint:
            state["optimizer"] = engine.optimizer.state_dict()
        uc = getattr(engine.module, "universal_checkpoint_info", None)
        if uc is not None:
            state[UNIVERSAL_CHECKPOINT_INFO] = uc() if callable(uc) else uc
        state.update(client_state or {})
        if engine.zero_optimization_partition_weights() or ce.is_data_parallel_writer(engine.seq_dp_rank):
            ce.save(state, get_ckpt_name(engine, save_dir, tag))

    if zero_sd is not None:
        zero_state = {OPTIMIZER_STATE_DICT: zero_sd, "ds_config": engine.config, "ds_version": __version__}
        # this rank's own shard: never shared between data-parallel writers
        kw = {"replicated": False} if getattr(ce, "supports_shared_write", False) else {}
        ce.save(zero_state, get_zero_ckpt_name(engine, save_dir, tag), **kw)
        if engine.global_rank == 0:
            _copy_recovery_script(os.path.join(save_dir, tag))

    if ce.is_decoupled():
        # every rank's writer process confirms its own files; only rank 0's also moves the ``latest`` marker
        if engine.global_rank == 0:
            ce.commit(info)
        else:
            ce.commit(type(info)(info.tag, info.save_dir, False))
    else:
        dist.barrier()
        ce.commit(info)
        if save_latest and engine.global_rank == 0:
            with open(os.path.join(save_dir, "latest"), "w") as fd:
                fd.write(tag)
        dist.barrier()
    return True


def _copy_recovery_script(dst_dir):
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "zero_to_fp32.py")
    try:
        shutil.copyfile(src, os.path.join(dst_dir, "zero_to_fp32.py"))
        os.chmod(os.path.join(dst_dir, "zero_to_fp32.py"), 0o755)
    except OSError as e:
        logger.warning(f"could not copy zero_to_fp32.py into the checkpoint: {e}")


# ---------------------------------------------------------------------------------------------------------
# load
# ---------------------------------------------------------------------------------------------------------
def load_engine_checkpoint(engine, load_dir, tag=None, load_module_strict=True, load_optimizer_states=True,
                           load_lr_scheduler_states=True, load_module_only=False, custom_load_fn=None):
    universal = engine.load_universal_checkpoint()
    if tag is None:
        latest = os.path.join(load_dir, "latest_universal" if universal else "latest")
        if os.path.isfile(latest):
            with open(latest) as fd:
                tag = fd.read().strip()
        else:
            if universal:
                raise ValueError(f"Invalid for universal checkpoint: {latest} does not exist")
            logger.warning(f"Unable to find latest file at {latest}, if trying to load latest checkpoint please "
                           "ensure this file exists or pass an explicit checkpoint tag when loading a checkpoint.")
            return None, None
    ce = engine.checkpoint_engine
    if universal:
        from .universal import load_universal_into_engine
        return load_universal_into_engine(engine, load_dir, tag, load_optimizer_states, load_lr_scheduler_states)

    path = get_ckpt_name(engine, load_dir, tag)
    if not os.path.exists(path):
        # stage mismatch tolerant lookup: try the other naming
        alt = os.path.join(load_dir, str(tag), f"mp_rank_{_mp_rank(engine):02d}_model_states.pt")
        alt3 = os.path.join(load_dir, str(tag), f"zero_pp_rank_{engine.seq_dp_rank}_mp_rank_{_mp_rank(engine):02d}_model_states.pt")
        path = alt if os.path.exists(alt) else alt3
    if not os.path.exists(path):
        from ..io import find_striped_writers
        if find_striped_writers(get_ckpt_name(engine, load_dir, tag)) is not None:
            path = get_ckpt_name(engine, load_dir, tag)  # written in stripes by several data-parallel ranks
        else:
            logger.warning(f"Client provided checkpoint load path: {path} does not exist")
            return None, None
    if not os.path.exists(path):
        from ..io import StripedReader
        with StripedReader(path) as f:
            ckpt = torch.load(f, map_location="cpu", weights_only=False)
    else:
        ckpt = ce.load(path, map_location="cpu")
    zero3 = engine.zero_optimization_partition_weights()

    have_zero_files = engine.optimizer is not None and os.path.exists(get_zero_ckpt_name(engine, load_dir, tag))
    if not zero3:
        if engine.has_moe_layers:
            from ..moe.checkpoint import load_moe_experts
            load_moe_experts(engine, load_dir, tag, ckpt["module"])
        engine.load_module_state_dict(ckpt, strict=load_module_strict, custom_load_fn=custom_load_fn)
    else:
        # buffers + persistent params come from the model_states file; sharded params from fp32 shards
        sd = ckpt["module"]
        bufs = dict(engine.module.named_buffers())
        for name, val in sd.items():
            if name in bufs and bufs[name] is not None and val.numel() == bufs[name].numel():
                bufs[name].data.copy_(val.to(bufs[name].device))
        if engine.optimizer is not None and hasattr(engine.optimizer, "units") and not have_zero_files:
            _zero3_load_persistent_from_module(engine, ckpt)

    engine.loaded_checkpoint_dp_world_size = ckpt.get("dp_world_size")
    engine.loaded_checkpoint_mp_world_size = ckpt.get("mp_world_size")
    if not load_module_only:
        engine.global_steps = ckpt.get("global_steps", 0)
        engine.global_samples = ckpt.get("global_samples", engine.global_steps * engine.train_batch_size())
        engine.skipped_steps = ckpt.get("skipped_steps", 0)
        if load_lr_scheduler_states and engine.lr_scheduler is not None and ckpt.get("lr_scheduler") is not None:
            engine.lr_scheduler.load_state_dict(ckpt["lr_scheduler"])
        if getattr(engine, "random_ltd_scheduler", None) is not None and ckpt.get("random_ltd") is not None:
            engine.random_ltd_scheduler.load_state_dict(ckpt["random_ltd"])
        sampler = engine.get_data_sampler() if hasattr(engine, "get_data_sampler") else None
        if sampler is not None and hasattr(sampler, "load_state_dict") and ckpt.get("data_sampler") is not None:
            sampler.load_state_dict(ckpt["data_sampler"])

    if engine.optimizer is not None:
        if have_zero_files:
            zsd = ce.load(get_zero_ckpt_name(engine, load_dir, tag), map_location="cpu")[OPTIMIZER_STATE_DICT]
            _load_zero_state(engine, zsd, load_optimizer_states and not load_module_only,
                             load_from_fp32=engine.zero_load_from_fp32_weights() or zero3)
        elif load_optimizer_states and not load_module_only:
            osd = None
            if engine.has_moe_layers:
                from ..moe.checkpoint import load_expert_optimizer_state
                osd = load_expert_optimizer_state(engine, load_dir, tag)
            osd = osd if osd is not None else ckpt.get("optimizer")
            if osd is not None:
                engine.optimizer.load_state_dict(osd)
            engine.optimizer.refresh_lp_from_params() if not zero3 else None
        else:
            engine.optimizer.refresh_lp_from_params()
        if zero3 and hasattr(engine.optimizer, "units"):
            _zero3_load_frozen(engine, ckpt)

    reserved = {"module", "buffer_names", "optimizer", "param_shapes", "frozen_param_shapes", "frozen_param_fragments",
                "shared_params", "lr_scheduler", "data_sampler", "random_ltd", "sparse_tensor_module_names",
                "skipped_steps", "global_steps", "global_samples", "dp_world_size", "mp_world_size",
                "checkpoint_parallel_dimensions", "ds_config", "ds_version", UNIVERSAL_CHECKPOINT_INFO}
    client_state = {k: v for k, v in ckpt.items() if k not in reserved}
    return os.path.join(load_dir, str(tag)), client_state


def _matching_expert_units(z, zsd):
    """(unit, saved entry) for the expert units whose saved shard fits the current topology."""
    saved = zsd.get("expert_units") or {}
    out = []
    for u in z.units:
        if getattr(u, "group_key", "dp") == "dp":
            continue
        e = saved.get(u.name)
        if e is None:
            continue
        if list(e["names"]) != list(u.names) or e["world"] != u.world or e["rank"] != u.rank or \
                e["fp32"].numel() != u.master_shard.numel():
            logger.warning(f"expert unit {u.name}: saved shard (world {e['world']}, rank {e['rank']}) does not match the "
                           f"current topology (world {u.world}, rank {u.rank}); its optimizer state is not restored")
            continue
        out.append((u, e))
    return out


def _load_zero_state(engine, zsd, load_optimizer_states, load_from_fp32=True):
    z = engine.optimizer
    if hasattr(z, "_join_step"):
        z._join_step()
    saved_world = zsd.get("partition_count", z._shard_world)
    if isinstance(saved_world, (list, tuple)):
        saved_world = max(saved_world)
    if saved_world != z._shard_world:
        raise RuntimeError(f"ZeRO checkpoint partition_count={saved_world} differs from the current shard world "
                           f"{z._shard_world}. Convert with ds_to_universal and load with checkpoint.load_universal.")
    lay = ReferenceLayout(z)
    stage_saved = zsd.get("zero_stage", z.stage)
    if (stage_saved >= 3) != (z.stage >= 3):
        raise RuntimeError(f"checkpoint zero_stage={stage_saved} is not loadable into stage {z.stage} directly; "
                           "use the universal checkpoint path")
    z.loss_scaler = zsd.get("loss_scaler", z.loss_scaler)
    z.dynamic_loss_scale = zsd.get("dynamic_loss_scale", z.dynamic_loss_scale)
    z.overflow = zsd.get("overflow", False)
    fp32 = zsd[FP32_FLAT_GROUPS] if z.stage >= 3 else zsd[SINGLE_PARTITION_OF_FP32_GROUPS]
    fp32 = [t.float() for t in fp32]
    if z.stage >= 3 and len(fp32) != len(lay.group_params):
        fp32 = _regroup_stage3(fp32, lay)
    if load_from_fp32:

        n_opt_groups = len(z.optimizer.param_groups)

        def write_master(u, shard):
            # only the regions of optimizer-managed parameters: frozen parameters are not in the fp32 flats
            for run in u.runs_in_shard():
                if run.group_idx < n_opt_groups:
                    u.master_shard[run.start:run.end].copy_(shard[run.start:run.end].to(u.master_shard.dtype))

        lay.import_(fp32, write_master)
        for u, entry in _matching_expert_units(z, zsd):
            u.master_shard.copy_(entry["fp32"].to(u.master_shard.dtype))
        z._lp_written_by_step = False
        z._sync_lp_from_master()
        z._post_step_gather()
    else:
        z.refresh_lp_from_params()
    if load_optimizer_states:
        base = zsd[OPTIMIZER_STATE_DICT] if z.stage >= 3 else zsd[BASE_OPTIMIZER_STATE]
        states = base["state"]
        for g, saved in zip(z.optimizer.param_groups, base.get("param_groups", [])):
            g.update({k: v for k, v in saved.items() if k != "params"})
        n_groups = len(lay.group_params)
        tensor_keys = sorted({k for gi in range(n_groups) for k, v in states.get(gi, {}).items()
                              if torch.is_tensor(v) and v.numel() > 1})
        if z.stage >= 3 and len(states) != n_groups:
            states = _regroup_stage3_states(states, lay, tensor_keys)
        for k in tensor_keys:
            parts = [states[gi][k].float() if k in states.get(gi, {}) else torch.zeros(lay.partition_size[gi])
                     for gi in range(n_groups)]
            shards = {}
            lay.import_(parts, lambda u, s: shards.__setitem__(u.uid, s.clone()))
            for (skey, a, b, gi, u, v) in z._segments:
                if u.uid not in shards:
                    continue  # expert units: restored from their per-unit entries below
                src = shards[u.uid][a - u.shard_offset:b - u.shard_offset]
                z.optimizer.state[v][k] = src.to(device=v.device, dtype=_state_dtype(z, v)).clone()
        expert_state = {u.uid: entry["state"] for u, entry in _matching_expert_units(z, zsd)}
        for (skey, a, b, gi, u, v) in z._segments:
            for k, shard in expert_state.get(u.uid, {}).items():
                z.optimizer.state[v][k] = shard[a - u.shard_offset:b - u.shard_offset].to(
                    device=v.device, dtype=_state_dtype(z, v)).clone()
        for (skey, a, b, gi, u, v) in z._segments:
            for k, val in states.get(gi, {}).items():
                if not (torch.is_tensor(val) and val.numel() > 1):
                    z.optimizer.state[v][k] = val
        if hasattr(z.optimizer, "on_state_loaded"):
            z.optimizer.on_state_loaded()
    if getattr(z, "coordinator", None) is not None:
        z.coordinator.release_all()  # units a plan keeps gathered hold pre-load contents


def _state_dtype(z, v):
    return v.dtype


def _regroup_stage3(flat_list, lay):
    """Reference stage-3 checkpoints hold one flat per *sub-group*; concatenate and re-split per param group."""
    cat = torch.cat([t.reshape(-1) for t in flat_list])
    out, off = [], 0
    for gi in range(len(lay.group_params)):
        n = lay.partition_size[gi]
        out.append(cat[off:off + n])
        off += n
    return out


def _regroup_stage3_states(states, lay, tensor_keys):
    keys = sorted(states.keys())
    out = {}
    for k in tensor_keys:
        cat = torch.cat([states[i][k].reshape(-1) for i in keys if k in states[i]])
        off = 0
        for gi in range(len(lay.group_params)):
            n = lay.partition_size[gi]
            out.setdefault(gi, {})[k] = cat[off:off + n]
            off += n
    for gi in range(len(lay.group_params)):
        for kk, vv in states[keys[0]].items():
            if not (torch.is_tensor(vv) and vv.numel() > 1):
                out.setdefault(gi, {})[kk] = vv
    return out
