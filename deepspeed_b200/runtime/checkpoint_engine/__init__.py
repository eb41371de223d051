"""Checkpoint engines: torch (sync), fast (pinned double-buffer native writer), decoupled (writer process).

Interface follows reference ``runtime/checkpoint_engine/checkpoint_engine.py:21-63``
(``create/save/load/commit/makedirs/is_data_parallel_writer/is_decoupled``); factory ``utils.py:15``.
"""
import multiprocessing as mp
from enum import Enum
import os
import queue as _q

import torch

from ...utils.logging import log_dist, logger


class CheckpointCommitInfo:

    def __init__(self, tag, save_dir, save_latest):
        self.tag, self.save_dir, self.save_latest = tag, save_dir, save_latest


class CheckpointEngine:

    def __init__(self, config_params=None):
        self.name = None
        self.config_params = config_params

    def create(self, info):
        log_dist(f"[{type(self).__name__}] Checkpoint {getattr(info, 'tag', info)} is about to be saved!", [0])

    def save(self, state_dict, path):
        raise NotImplementedError

    def load(self, path, map_location=None):
        return torch.load(path, map_location=map_location, weights_only=False)

    def commit(self, info):
        return True

    def makedirs(self, path, exist_ok=False):
        os.makedirs(path, exist_ok=exist_ok)

    def is_data_parallel_writer(self, dp_rank):
        return dp_rank == 0

    def is_decoupled(self):
        return False

    def set_commit_info(self, info):
        self._info = info

    def get_commit_info(self):
        return getattr(self, "_info", None)

    def cleanup(self):
        pass

    def preserves_storage_sharing(self):
        return True


class TorchCheckpointEngine(CheckpointEngine):

    def save(self, state_dict, path):
        torch.save(state_dict, path)

    def commit(self, info):
        logger.info(f"[Torch] Checkpoint {getattr(info, 'tag', info)} is ready now!")
        return True


class FastCheckpointEngine(CheckpointEngine):
    """Serialises with torch's pickle protocol into a pinned double-buffered native writer
    (:class:`deepspeed_b200.io.FastFileWriter`, O_DIRECT + worker threads).

    ``dp_writer_config`` (``runtime/model_checkpointing``): with a ``data_parallel`` unit in the writer section, the
    ranks elected among those holding the same state each write a stripe set of the model-state file
    (``<file>-<rank>.<N>``); ranks that were not elected write no replicated state at all. ``load`` accepts either
    layout. ``optimize_dp_state`` is off when the weights themselves are partitioned (ZeRO-3): nothing is replicated."""

    supports_shared_write = True

    def __init__(self, config_params=None, dp_writer_config="rank0", optimize_dp_state=True):
        super().__init__(config_params)
        from ..model_checkpointing import CheckpointWriterFactory, get_checkpoint_writer_config
        self.name = "FastCheckpointEngine"
        w = getattr(config_params, "writer", None)
        wd = w.model_dump() if hasattr(w, "model_dump") else dict(w or {})
        wd["type"] = "fast"
        if not wd.get("io_buffer_size"):
            wd["io_buffer_size"] = 1 << 28
        self._writer_cfg = get_checkpoint_writer_config({"writer": wd})
        self.dp_writer_config = dp_writer_config
        self.optimize_dp_state = optimize_dp_state
        self.io_buffer_size = self._writer_cfg["io_buffer_size"]
        self.double = self._writer_cfg["io_buffer_double"]
        self._factory = None
        self._factory_cls = CheckpointWriterFactory
        self._aio = getattr(config_params, "aio_config", None)

    def _shares(self):
        dp = self.dp_writer_config
        return self.optimize_dp_state and dp is not None and dp != "rank0" and dp.world_size > 1

    def save(self, state_dict, path, replicated=True):
        """``replicated``: the state is identical on the ranks of the data-parallel group (model states); per-rank
        files (ZeRO shards) are always written whole."""
        if self._factory is None:
            dp = None if self.dp_writer_config == "rank0" else self.dp_writer_config
            self._factory = self._factory_cls(self._writer_cfg, self._aio, dp)
        writer = self._factory.create_writer(path, bool(replicated and self._shares()))
        try:
            torch.save(state_dict, writer, _use_new_zipfile_serialization=False)
        finally:
            self._factory.release_writer()

    def load(self, path, map_location=None):
        if not os.path.exists(path):
            from ...io import StripedReader, find_striped_writers
            n = find_striped_writers(path)
            if n is not None:
                with StripedReader(path, n) as f:
                    return torch.load(f, map_location=map_location, weights_only=False)
        return torch.load(path, map_location=map_location, weights_only=False)

    def is_data_parallel_writer(self, dp_rank):
        if self.dp_writer_config == "rank0":
            return dp_rank == 0
        return self.dp_writer_config is not None


class DecoupledEvent(Enum):
    SAVE_EVENT = 1
    COMMIT_EVENT = 2
    EXIT_EVENT = 3


class CheckpointSize:
    """Bytes a checkpoint added to its directory (sizes before / after), in GiB."""

    def __init__(self):
        self._pre = self._post = self._gigabytes = None

    def gb_size(self):
        return self._gigabytes

    def set_pre_size(self, size):
        self._pre = size

    def set_post_size(self, size):
        self._post = size
        self._gigabytes = (self._post - self._pre) / (1024**3)


def _dir_bytes(path):
    total = 0
    for root, _, files in os.walk(path):
        for f in files:
            try:
                total += os.path.getsize(os.path.join(root, f))
            except OSError:
                pass
    return total


class _WriterSection:  # picklable stand-in for the pydantic ``checkpoint`` section inside the writer process

    def __init__(self, writer):
        self.writer = writer


def _writer_main(q, done, fast=None):
    """Writer process. ``fast``: (writer section as a dict, data-parallel writer config, optimize_dp_state) - saves go
    through :class:`FastCheckpointEngine` (pinned staging + native async writes, stripes shared with the other
    data-parallel writers); without it plain ``torch.save``."""
    engine = None
    if fast is not None:
        try:
            engine = FastCheckpointEngine(_WriterSection(fast[0]), fast[1], fast[2])
        except Exception as e:  # noqa: BLE001
            print(f"[Decoupled] fast writer unavailable in the writer process ({e}); using torch.save", flush=True)
    size = CheckpointSize()
    while True:
        item = q.get()
        if item is None:
            break
        kind, payload = item
        if kind in ("save", DecoupledEvent.SAVE_EVENT):
            sd, path, replicated = (tuple(payload) + (True, ))[:3]
            if size._pre is None:
                size.set_pre_size(_dir_bytes(os.path.dirname(path)))
            if engine is not None:
                engine.save(sd, path, replicated=replicated)
            else:
                torch.save(sd, path)
        elif kind in ("commit", DecoupledEvent.COMMIT_EVENT):
            save_dir, tag, save_latest = payload
            if save_latest:
                with open(os.path.join(save_dir, "latest"), "w") as f:
                    f.write(tag)
            size.set_post_size(_dir_bytes(os.path.join(save_dir, str(tag))))
            size._pre = None
            done.put(tag)


class DecoupledCheckpointEngine(CheckpointEngine):
    """Hands CPU copies of the state to a separate writer *process*; commit is confirmed at the next GAS
    boundary (reference ``decoupled_checkpoint_engine.py``; engine.py:3452-3453). With a ``fast`` writer section the
    process writes through :class:`FastCheckpointEngine`, including the data-parallel shared write."""
    supports_shared_write = True

    def __init__(self, config_params=None, dp_writer_config="rank0", optimize_dp_state=True):
        super().__init__(config_params)
        self.dp_writer_config = dp_writer_config
        w = getattr(config_params, "writer", None)
        wd = w.model_dump() if hasattr(w, "model_dump") else dict(w or {})
        fast = (wd, dp_writer_config, optimize_dp_state) if str(wd.get("type", "")).lower() == "fast" else None
        ctx = mp.get_context("spawn")
        self.q, self.done = ctx.Queue(), ctx.Queue()
        self.proc = ctx.Process(target=_writer_main, args=(self.q, self.done, fast), daemon=True)
        self.proc.start()
        self.pending = []

    def is_data_parallel_writer(self, dp_rank):
        if self.dp_writer_config == "rank0":
            return dp_rank == 0
        return self.dp_writer_config is not None

    def is_decoupled(self):
        return True

    def _alive(self):
        if not self.proc.is_alive():
            raise RuntimeError("decoupled checkpoint writer process died")

    @staticmethod
    def _to_cpu(obj):
        if torch.is_tensor(obj):
            return obj.detach().cpu().clone()
        if isinstance(obj, dict):
            return type(obj)((k, DecoupledCheckpointEngine._to_cpu(v)) for k, v in obj.items())
        if isinstance(obj, (list, tuple)):
            return type(obj)(DecoupledCheckpointEngine._to_cpu(v) for v in obj)
        return obj

    def save(self, state_dict, path, replicated=True):
        self._alive()
        self.q.put(("save", (self._to_cpu(state_dict), path, replicated)))

    def commit(self, info):
        self._alive()
        self.q.put(("commit", (info.save_dir, info.tag, info.save_latest)))
        self.pending.append(info.tag)
        return True

    def commit_pending(self, block=False):
        while self.pending:
            try:
                tag = self.done.get(block=block, timeout=600 if block else None)
            except _q.Empty:
                return
            if tag in self.pending:
                self.pending.remove(tag)
                logger.info(f"[Decoupled] Checkpoint {tag} is ready now!")

    def cleanup(self):
        try:
            self.commit_pending(block=True)
            self.q.put(None)
            self.proc.join(timeout=30)
        except Exception:
            pass


class NebulaCheckpointEngine(CheckpointEngine):
    """Azure Nebula service (reference ``nebula_checkpoint_engine.py``): saves go to the service's tiered storage under the
    checkpoint tag, a dummy file marks each partition on disk so that the usual file discovery keeps working. Needs the
    external ``torch_nebula`` package."""

    def __init__(self, config_params=None):
        super().__init__(config_params)
        try:
            import torch_nebula
        except ImportError as e:
            raise RuntimeError("the nebula checkpoint engine needs the torch_nebula package (Azure ML images)") from e
        self.nb = torch_nebula
        self.name = "NebulaCheckpointEngine"
        self.checkpoint, self.tag_flag = None, None
        self.enable_nebula_load = config_params.enable_nebula_load
        self.nebula_load_path = config_params.load_path or config_params.persistent_storage_path
        self.nb.init(persistent_storage_path=config_params.persistent_storage_path,
                     persistent_time_interval=config_params.persistent_time_interval,
                     num_of_version_in_retention=config_params.num_of_version_in_retention)

    @staticmethod
    def _tag_of(path):
        return os.path.basename(os.path.dirname(path))

    def create(self, info):
        log_dist(f"[Nebula] Start Checkpoint for tag:{info.tag}", [0])
        self.checkpoint = self.nb.Checkpoint(info.tag, -2)  # -2: complete only on an explicit commit

    def save(self, state_dict, path):
        torch.save("", path)  # placeholder so that tag / file discovery works
        self.checkpoint.save(os.path.basename(path), state_dict)

    def load(self, path, map_location=None):
        tag = self._tag_of(path)
        if not self.enable_nebula_load and (self.tag_flag is None or self.tag_flag == tag):
            self.tag_flag = tag
            return torch.load(path, map_location=map_location, weights_only=False)
        ck = self.nb.get_latest_checkpoint(persist_path=self.nebula_load_path) \
            if tag in (None, "latest", "latest_universal") else \
            self.nb.get_checkpoint(tag=tag, persist_path=self.nebula_load_path)
        if ck is None or ck.tag == "":
            ck = self.nb.get_latest_checkpoint(persist_path=self.nebula_load_path)
        if ck is None or ck.tag == "":
            logger.info(f"[Nebula] no checkpoint for tag {tag}; falling back to the file {path}")
            return torch.load(path, map_location=map_location, weights_only=False)
        self.tag_flag = ck.tag
        return ck.load(os.path.basename(path), map_location=map_location)

    def commit(self, info):
        tag = getattr(info, "tag", info)
        log_dist(f"[Nebula] all files for {tag} are saved in tier1; persisting to tier3 asynchronously", [0])
        return self.checkpoint.commit()


class DataStatesCheckpointEngine(CheckpointEngine):
    """DataStates-LLM asynchronous multi-level checkpointing (reference ``datastates_checkpoint_engine.py``). Needs the
    external ``datastates`` package."""

    def __init__(self, deepspeed_config, rank):
        super().__init__(deepspeed_config)
        try:
            from datastates import CheckpointEngine as _Engine
        except ImportError as e:
            raise RuntimeError("Please install DataStates from https://github.com/DataStates/datastates-llm.") from e
        self.ckpt_engine = _Engine(deepspeed_config, rank)
        self.commit_info = None

    def create(self, info):
        self.commit_info = info

    def save(self, state_dict, path):
        return self.ckpt_engine.save(state_dict, path)

    def load(self, path, map_location=None):
        return self.ckpt_engine.load(path, map_location)

    def commit(self, info):
        assert info == self.commit_info
        self.ckpt_engine.wait(persist=True)
        self.commit_info = None
        return True

    def cleanup(self):
        if getattr(self, "ckpt_engine", None) is not None:
            self.commit(self.commit_info)
            del self.ckpt_engine
            self.ckpt_engine = None

    def is_decoupled(self):
        return False

    def preserves_storage_sharing(self):
        return False


def create_checkpoint_engine(config, zero_stage=0, has_moe_layers=False):
    neb = getattr(config, "nebula_config", None) or {}
    if isinstance(neb, dict) and neb.get("enabled", False):
        from ...nebula.config import DeepSpeedNebulaConfig
        try:
            return NebulaCheckpointEngine(DeepSpeedNebulaConfig({"nebula": neb}))
        except RuntimeError as e:
            logger.warning(f"{e}; falling back to the torch checkpoint engine")
    dst = getattr(config, "datastates_config", None) or {}
    if isinstance(dst, dict) and dst.get("enabled", False):
        try:
            from ... import comm as _dist
            return DataStatesCheckpointEngine(dst, _dist.get_rank() if _dist.is_initialized() else 0)
        except RuntimeError as e:
            logger.warning(f"{e}; falling back to the torch checkpoint engine")
    ck = getattr(config, "checkpoint_config", None)
    w = getattr(ck, "writer", None) if ck is not None else None
    kind = (getattr(w, "type", "torch") or "torch").lower() if w is not None else "torch"
    if w is not None and getattr(w, "decoupled", False):
        kind = "decoupled"
    if kind == "fast":
        try:
            dp_cfg = "rank0"
            unit = getattr(w, "data_parallel", None)
            from ... import comm as _dist
            if unit and _dist.is_initialized():
                from ...utils import groups as _groups
                from ..model_checkpointing import create_data_parallel_writer_config
                dp_cfg = create_data_parallel_writer_config(_groups, str(unit).upper(), zero_stage, has_moe_layers)
            return FastCheckpointEngine(ck, dp_cfg, optimize_dp_state=int(zero_stage) != 3)
        except Exception as e:
            logger.warning(f"FastCheckpointEngine unavailable ({e}); falling back to torch")
            return TorchCheckpointEngine(ck)
    if kind == "decoupled":
        dp_cfg = "rank0"
        unit = getattr(w, "data_parallel", None)
        from ... import comm as _dist
        if unit and _dist.is_initialized() and (getattr(w, "type", "") or "").lower() == "fast":
            from ...utils import groups as _groups
            from ..model_checkpointing import create_data_parallel_writer_config
            dp_cfg = create_data_parallel_writer_config(_groups, str(unit).upper(), zero_stage, has_moe_layers)
        return DecoupledCheckpointEngine(ck, dp_cfg, optimize_dp_state=int(zero_stage) != 3)
    return TorchCheckpointEngine(ck)
