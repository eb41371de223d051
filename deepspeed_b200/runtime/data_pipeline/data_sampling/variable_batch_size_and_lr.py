"""Dynamic batching by token budget + per-batch LR scaling (reference
``data_sampling/variable_batch_size_and_lr.py``). Micro-batches hold as many sequences as fit in ``max_tokens``; the
LR of every step is rescaled (linear / sqrt) by that step's effective batch size relative to the base batch size."""
import math
import os
import pathlib
import random

import numpy as np
from torch.optim.lr_scheduler import LRScheduler
from torch.optim.optimizer import Optimizer
from torch.utils.data import DataLoader, DistributedSampler

from ....utils.logging import logger
from .data_analyzer import DistributedDataAnalyzer
from .indexed_dataset import MMapIndexedDataset


def batch_by_seqlens(seqlens, max_tokens, sequence_ids_per_mb=None, min_batch_size=1, max_batch_size=None,
                     sequence_picking_order="dataloader", effective_batch_size=1,
                     required_microbatches_of_same_size=False, verbose=False, seed=None):
    """-> (microbatch_ids [(batch_id, [sample ids])...], batch_sizes, batch_max_seqlens).

    Samples are taken in ``sequence_picking_order`` and packed greedily: a group of ``k`` interleaved micro-batches
    (k = effective_batch_size when equal sizes are required, else 1) grows while every micro-batch stays within
    ``max_tokens`` / ``max_batch_size``. Growth stops at the first violation (limits are monotone), which makes the
    packing linear in the dataset size."""
    assert sequence_picking_order in ("random", "seqlen", "dataloader")
    ids = range(len(seqlens)) if sequence_ids_per_mb is None else sequence_ids_per_mb
    metrics = [(seqlens[i], i) for i in ids]
    if sequence_picking_order == "random":
        random.Random(seed).shuffle(metrics)
    elif sequence_picking_order == "seqlen":
        metrics.sort()
    too_long = {i for v, i in metrics if v > max_tokens}
    if too_long:
        logger.warning(f"{len(too_long)} samples ignored: their metric exceeds max_tokens={max_tokens}")
        metrics = [m for m in metrics if m[1] not in too_long]

    k = effective_batch_size if required_microbatches_of_same_size else 1
    microbatches, pos, n = [], 0, len(metrics)
    while pos < n:
        tokens, counts, best = [0] * k, [0] * k, 0
        taken = 0
        while pos + taken + k <= n:
            # try to add one more sequence to each of the k micro-batches
            ok = True
            for j in range(k):
                v = metrics[pos + taken + j][0]
                if tokens[j] + v > max_tokens or (max_batch_size and counts[j] + 1 > max_batch_size):
                    ok = False
                    break
            if not ok:
                break
            for j in range(k):
                tokens[j] += metrics[pos + taken + j][0]
                counts[j] += 1
            taken += k
            if not min_batch_size or counts[0] >= min_batch_size:
                best = taken
        if best == 0:
            break
        group = metrics[pos:pos + best]
        microbatches += [group[j::k] for j in range(k)]
        pos += best
    assert len(microbatches) >= effective_batch_size, "not enough datapoints to create a single sample per dataloader"
    microbatches = microbatches[:len(microbatches) - len(microbatches) % effective_batch_size]

    microbatch_ids, batch_sizes, batch_max_seqlens = [], [], []
    for first in range(0, len(microbatches), effective_batch_size):
        batch_id = first // effective_batch_size
        mbs = microbatches[first:first + effective_batch_size]
        batch_sizes.append(sum(len(mb) for mb in mbs))
        batch_max_seqlens.append(max(v for mb in mbs for v, _ in mb))
        microbatch_ids += [(batch_id, [i for _, i in mb]) for mb in mbs]
        if verbose:
            logger.info(f"batch {batch_id}: {len(mbs)} microbatches, {batch_sizes[-1]} sequences, "
                        f"tokens per microbatch {[sum(v for v, _ in mb) for mb in mbs]}")
    return microbatch_ids, batch_sizes, batch_max_seqlens


def scale_lr(base_batch_size, batch_size, base_lr=1, method="linear"):
    if method == "linear":
        return base_lr * batch_size / base_batch_size
    if method == "sqrt":
        return base_lr * math.sqrt(batch_size / base_batch_size)
    if method is None or str(method).upper() == "NONE":
        return base_lr
    raise ValueError(f"Unknown scaling method: {method}")


def dataloader_for_variable_batch_size(dataset, microbatch_ids, batch_max_seqlens, dataloader_rank=0,
                                       dataloader_batch_size=1, dataloader_num_replicas=1, dataloader_collate_fn=None,
                                       dataloader_num_workers=2, dataloader_pin_memory=False,
                                       required_microbatches_of_same_seqlen=False, sample_padding_fn=None):
    """The "dataset" handed to the DataLoader is the list of micro-batches; the collate function materialises them."""
    sampler = DistributedSampler(dataset=microbatch_ids, num_replicas=dataloader_num_replicas, rank=dataloader_rank,
                                 shuffle=False, drop_last=False)

    def collate_fn_wrapper(list_microbatch_ids):
        batch = []
        for batch_id, sample_ids in list_microbatch_ids:
            rows = [dataset[i] for i in sample_ids]
            if required_microbatches_of_same_seqlen:
                assert sample_padding_fn is not None, "sample_padding_fn is required to pad to a common seqlen"
                width = batch_max_seqlens[batch_id]
                assert all(len(r) <= width for r in rows)
                rows = [sample_padding_fn(r, width) for r in rows]
            batch += rows
        return dataloader_collate_fn(batch) if dataloader_collate_fn else batch

    loader = DataLoader(dataset=microbatch_ids, batch_size=dataloader_batch_size, sampler=sampler,
                        num_workers=dataloader_num_workers, collate_fn=collate_fn_wrapper,
                        pin_memory=dataloader_pin_memory)
    io_kwargs = dict(dataset=microbatch_ids, batch_size=dataloader_batch_size, pin_memory=dataloader_pin_memory,
                     data_sampler=sampler, collate_fn=collate_fn_wrapper, num_local_io_workers=dataloader_num_workers)
    return loader, io_kwargs


class VariableBatchSizeLR(LRScheduler):
    """Wraps a scheduler; after each of its steps multiplies the LR by the batch-size ratio of the coming batch."""

    @property
    def optimizer(self):
        return self.base_lr_scheduler.optimizer

    def __init__(self, lr_scheduler, base_batch_size, batch_sizes, dataloader, lr_scaling_method="linear",
                 last_epoch=-1, verbose=False):
        self.batch_sizes, self.base_batch_size = batch_sizes, base_batch_size
        self.lr_scaling_method, self.dataloader = lr_scaling_method, dataloader
        self.base_lr_scheduler = lr_scheduler
        self.base_lrs = self.base_lr_scheduler.get_lr()
        self.last_epoch, self.verbose = last_epoch, verbose
        self.step(0)

    _KEYS = ("base_batch_size", "lr_scaling_method", "batch_sizes", "base_lrs", "last_epoch", "verbose")

    def state_dict(self):
        return {"base_lr_scheduler": self.base_lr_scheduler.state_dict(), **{k: getattr(self, k) for k in self._KEYS}}

    def load_state_dict(self, state_dict):
        self.base_lr_scheduler.load_state_dict(state_dict["base_lr_scheduler"])
        for k in self._KEYS:
            setattr(self, k, state_dict[k])

    def get_last_lr(self):
        return self.base_lr_scheduler._last_lr

    def get_lr(self):
        return [g["lr"] for g in self.base_lr_scheduler.optimizer.param_groups]

    def step(self, epoch=None):
        unscaled = self.base_lrs if epoch == 0 else self.get_last_lr()
        for group, lr in zip(self.base_lr_scheduler.optimizer.param_groups, unscaled):
            group["lr"] = lr
        self.base_lr_scheduler.step(epoch)
        self.last_epoch = self.last_epoch + 1 if epoch is None else epoch
        batch_size = self.batch_sizes[self.last_epoch % len(self.batch_sizes)]
        for group in self.base_lr_scheduler.optimizer.param_groups:
            group["lr"] = scale_lr(self.base_batch_size, batch_size, group["lr"], self.lr_scaling_method)
        if self.verbose:
            logger.info(f"next batch {self.last_epoch}: batch size {batch_size}, lr {self.get_lr()}")


def lr_scheduler_for_variable_batch_size(base_batch_size, batch_sizes, dataloader, lr_scheduler_or_optimizer,
                                         lr_scaling_method="linear", verbose=False):

    class StubLRScheduler(LRScheduler):
        """Constant LR, taken from the optimizer's param groups."""

        def get_lr(self):
            return self.base_lrs

    if isinstance(lr_scheduler_or_optimizer, Optimizer):
        sched = StubLRScheduler(lr_scheduler_or_optimizer)
    elif hasattr(lr_scheduler_or_optimizer, "optimizer"):
        sched = lr_scheduler_or_optimizer
    else:
        raise ValueError(f"Unknown type for lr_scheduler_or_optimizer: {type(lr_scheduler_or_optimizer)}")
    return VariableBatchSizeLR(sched, base_batch_size, batch_sizes, dataloader, lr_scaling_method, verbose=verbose)


def get_dataloader_and_lr_scheduler_for_variable_batch_size(
        dataset, dataset_seqlens, max_tokens, effective_batch_size, dataset_filter_ids=None,
        lr_scaling_method="linear", min_batch_size=1, max_batch_size=None, sequence_picking_order="dataloader",
        dataloader_batch_size=1, dataloader_rank=0, dataloader_num_replicas=1, dataloader_num_workers=0,
        dataloader_collate_fn=None, dataloader_pin_memory=False, lr_scheduler_or_optimizer=None,
        required_microbatches_of_same_size=False, required_microbatches_of_same_seqlen=False, sample_padding_fn=None,
        verbose=False, seed=None):
    microbatch_ids, batch_sizes, batch_max_seqlens = batch_by_seqlens(
        dataset_seqlens, max_tokens, dataset_filter_ids, min_batch_size, max_batch_size, sequence_picking_order,
        effective_batch_size, required_microbatches_of_same_size, verbose, seed)
    loader, io_kwargs = dataloader_for_variable_batch_size(
        dataset, microbatch_ids, batch_max_seqlens, dataloader_rank, dataloader_batch_size, dataloader_num_replicas,
        dataloader_collate_fn, dataloader_num_workers, dataloader_pin_memory, required_microbatches_of_same_seqlen,
        sample_padding_fn)
    sched = lr_scheduler_for_variable_batch_size(effective_batch_size, batch_sizes, loader, lr_scheduler_or_optimizer,
                                                 lr_scaling_method, verbose)
    return loader, sched, io_kwargs


def get_dataloader_and_lr_scheduler_for_variable_batch_size_deepspeed(dataset, engine, dataset_seqlens=None,
                                                                      dataset_filter_ids=None,
                                                                      dataloader_collate_fn=None,
                                                                      sample_padding_fn=None, batch_seqlens_fn=None):
    """Engine-aware wrapper: sequence lengths come from the caller, from the analyzer's ``seqlen`` metric files under
    ``dynamic_batching.metrics_path``, or are produced by running the analyzer with ``batch_seqlens_fn``."""
    from ..config import get_data_efficiency_config
    from ...pipe.engine import PipelineEngine
    de = get_data_efficiency_config({"data_efficiency": engine._config.data_efficiency_config})
    ds_cfg = de["data_sampling"]
    bc = ds_cfg["dynamic_batching"]
    assert bc["enabled"], "Dynamic batching is not enabled in the config"
    if dataset_seqlens is None:
        metrics_path = bc["metrics_path"]
        path = os.path.join(metrics_path, "seqlen/seqlen_sample_to_metric")
        if not MMapIndexedDataset.exists(path):
            if batch_seqlens_fn is None:
                raise ValueError("batch_seqlens_fn must be provided if dataset_seqlens is not provided")
            DistributedDataAnalyzer(dataset=dataset, metric_functions=[batch_seqlens_fn],
                                    collate_fn=dataloader_collate_fn, batch_size=2**10,
                                    save_path=str(pathlib.Path(metrics_path)), metric_types=["single_value_per_sample"],
                                    metric_names=["seqlen"], device=str(engine.device).split(":")[0]).run_map_reduce()
        rows = MMapIndexedDataset(path, skip_warmup=True)
        assert len(rows) == len(dataset), "seqlen metric does not match the dataset; delete metrics_path and retry"
        dataset_seqlens = np.array([r[0] for r in rows], dtype=np.int64)
    is_pipe = isinstance(engine, PipelineEngine)
    from ....utils import groups
    _dp_rank = groups._get_data_parallel_rank
    return get_dataloader_and_lr_scheduler_for_variable_batch_size(
        dataset=dataset, dataset_filter_ids=dataset_filter_ids, dataset_seqlens=dataset_seqlens,
        effective_batch_size=engine.train_batch_size(), max_tokens=bc["max_tokens"],
        lr_scaling_method=bc["lr_scaling_method"], sequence_picking_order=bc["sequence_picking_order"],
        min_batch_size=bc["min_batch_size"], max_batch_size=bc["max_batch_size"],
        dataloader_batch_size=engine.train_micro_batch_size_per_gpu(), dataloader_rank=_dp_rank(),
        dataloader_num_replicas=engine.dp_world_size, dataloader_num_workers=ds_cfg["num_workers"],
        dataloader_collate_fn=dataloader_collate_fn, dataloader_pin_memory=ds_cfg["pin_memory"],
        sample_padding_fn=sample_padding_fn, lr_scheduler_or_optimizer=engine.lr_scheduler or engine.optimizer,
        required_microbatches_of_same_size=is_pipe, required_microbatches_of_same_seqlen=is_pipe,
        verbose=bc["verbose"], seed=de["seed"])
