"""Offline data analysis for curriculum learning (reference ``data_sampling/data_analyzer.py:23``): a map phase
computes per-sample metrics over a dataset split across workers x threads, a reduce phase produces

    <save>/<metric>/<metric>_sample_to_metric               row i   = metric value of sample i
    <save>/<metric>/<metric>_index_to_metric                row k   = k-th distinct metric value (ascending)
    <save>/<metric>/<metric>_index_to_sample                row k   = samples having that value
    <save>/<metric>/<metric>_index_to_sample_percentile_merged      ~100 rows of merged sample ids
    <save>/<metric>/<metric>_metric_value                   (``accumulate_value_over_samples`` metrics)

in the mmap indexed format the sampler consumes. The reduce is a single vectorised stable sort per metric instead of
the reference's per-value Python dictionaries."""
import os
from threading import Thread

import numpy as np
import torch
from torch.utils.data import BatchSampler, DataLoader, SequentialSampler, Subset

from .... import comm as dist
from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype, split_dataset

SINGLE = "single_value_per_sample"
ACCUM = "accumulate_value_over_samples"


def _write_rows(fname, rows, dtype):
    b = create_mmap_dataset_builder(fname, dtype)
    for r in rows:
        b.add_item_numpy(np.asarray(r))
    close_mmap_dataset_builder(b, fname)


def reduce_single_value_metric(values, metric_name, metric_save_path, total_num_samples=None):
    """values[i] = metric of sample i (int array). Writes the four index files and returns {value: count}."""
    values = np.asarray(values).astype(np.int64).reshape(-1)
    n = len(values) if total_num_samples is None else total_num_samples
    assert len(values) == n, "map results do not cover the dataset; did a map worker fail?"
    os.makedirs(metric_save_path, exist_ok=True)
    v_dtype = find_fit_int_dtype(int(values.min()), int(values.max()))
    s_dtype = find_fit_int_dtype(0, n - 1)
    base = f"{metric_save_path}/{metric_name}"
    _write_rows(f"{base}_sample_to_metric", values.reshape(-1, 1), v_dtype)
    order = np.argsort(values, kind="stable")
    uniq, starts, counts = np.unique(values[order], return_index=True, return_counts=True)
    groups = [order[s:s + c] for s, c in zip(starts, counts)]
    _write_rows(f"{base}_index_to_metric", uniq.reshape(-1, 1), v_dtype)
    _write_rows(f"{base}_index_to_sample", groups, s_dtype)
    step = max(1, len(groups) // 100)
    merged = [np.concatenate(groups[i:i + step], axis=None) for i in range(0, len(groups), step)]
    _write_rows(f"{base}_index_to_sample_percentile_merged", merged, s_dtype)
    return {int(u): int(c) for u, c in zip(uniq, counts)}


def metric_value_percentiles(num_sample_per_value, total_num_samples):
    """{percentile: smallest metric value covering that share of the samples}."""
    out, seen, pct = {}, 0, 1
    for value in sorted(num_sample_per_value):
        seen += num_sample_per_value[value]
        while pct <= 100 and seen >= total_num_samples * pct / 100.0:
            out[pct] = value
            pct += 1
    return out


class DataAnalyzer:

    def __init__(self, dataset, num_workers=1, worker_id=0, num_threads=1, num_threads_reduce=1, specific_threads=(),
                 batch_size=1, metric_names=(), metric_functions=(), metric_types=(), metric_dtypes=(),
                 save_path="./", collate_fn=None, custom_map_init=None, custom_map_update=None,
                 custom_map_finalize=None, custom_reduce=None, sample_indices=None):
        self.dataset, self.num_workers, self.worker_id = dataset, num_workers, worker_id
        self.num_threads, self.num_threads_reduce = num_threads, num_threads_reduce
        self.specific_threads, self.batch_size = list(specific_threads), batch_size
        self.metric_names, self.metric_functions = list(metric_names), list(metric_functions)
        self.metric_types, self.metric_dtypes = list(metric_types), list(metric_dtypes)
        self.save_path, self.collate_fn = save_path, collate_fn
        self.custom_map_init, self.custom_map_update = custom_map_init, custom_map_update
        self.custom_map_finalize, self.custom_reduce = custom_map_finalize, custom_reduce
        self.sample_indices = sample_indices

    def _partial(self, metric, worker, thread):
        return f"{self.save_path}/{metric}/worker{worker}_thread{thread}"

    # ---- map ---------------------------------------------------------------------------------
    def run_map_helper(self, thread_id):
        _, thread_splits = split_dataset(self.dataset, self.num_workers, self.worker_id, self.num_threads)
        start, end = (int(x) for x in thread_splits[thread_id])
        loader = DataLoader(Subset(self.dataset, range(start, end)),
                            batch_sampler=BatchSampler(SequentialSampler(range(end - start)), self.batch_size, False),
                            num_workers=0, collate_fn=self.collate_fn)
        if self.custom_map_init is not None:
            state = self.custom_map_init(thread_id, self.metric_names, self.metric_types, self.metric_dtypes,
                                         self.save_path, self.worker_id)
        else:
            state = [[] if t == SINGLE else None for t in self.metric_types]
        for data in loader:
            if self.custom_map_update is not None:
                self.custom_map_update(data, self.metric_types, self.metric_dtypes, self.metric_functions, state)
                continue
            for m, (fn, mtype) in enumerate(zip(self.metric_functions, self.metric_types)):
                vals = fn(data)
                vals = vals.cpu().numpy() if isinstance(vals, torch.Tensor) else np.asarray(vals)
                if mtype == SINGLE:
                    state[m].append(vals.reshape(-1))
                elif mtype == ACCUM:
                    state[m] = vals.astype(np.int64) if state[m] is None else state[m] + vals
                else:
                    raise ValueError(f"unknown metric type {mtype}")
        if self.custom_map_finalize is not None:
            self.custom_map_finalize(self.metric_types, self.metric_dtypes, state)
            return
        for m, (name, mtype) in enumerate(zip(self.metric_names, self.metric_types)):
            d = self._partial(name, self.worker_id, thread_id)
            os.makedirs(d, exist_ok=True)
            if mtype == SINGLE:
                vals = np.concatenate(state[m]) if state[m] else np.zeros(0, dtype=np.int64)
                np.save(f"{d}/{name}_sample_to_metric.npy", vals.astype(np.int64))
            else:
                np.save(f"{d}/{name}_metric_value.npy", np.asarray(state[m], dtype=np.int64))

    def run_map(self):
        threads = self.specific_threads or list(range(self.num_threads))
        if len(threads) > 1:
            ts = [Thread(target=self.run_map_helper, args=(t, )) for t in threads]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
        else:
            self.run_map_helper(threads[0])

    # ---- reduce ------------------------------------------------------------------------------
    def merge_map_results(self, dataset, metric_names, metric_types, save_path, num_workers, num_threads,
                          num_threads_reduce):
        total = len(dataset)
        for name, mtype in zip(metric_names, metric_types):
            parts = [self._partial(name, w, t) for w in range(num_workers) for t in range(num_threads)]
            mdir = f"{save_path}/{name}"
            if mtype == SINGLE:
                values = np.concatenate([np.load(f"{p}/{name}_sample_to_metric.npy") for p in parts])
                counts = reduce_single_value_metric(values, name, mdir, total)
                pct = metric_value_percentiles(counts, total)
                logger.info(f"metric {name}: range [{min(counts)}, {max(counts)}], median {pct.get(50)}")
            elif mtype == ACCUM:
                acc = sum(np.load(f"{p}/{name}_metric_value.npy") for p in parts)
                dt = find_fit_int_dtype(int(acc.min()), int(acc.max()))
                _write_rows(f"{mdir}/{name}_metric_value", [acc], dt)

    def run_reduce(self):
        fn = self.custom_reduce or self.merge_map_results
        fn(self.dataset, self.metric_names, self.metric_types, self.save_path, self.num_workers, self.num_threads,
           self.num_threads_reduce)

    def run_map_reduce(self, comm_group=None):
        self.run_map()
        if dist.is_initialized():
            dist.barrier(group=comm_group)
        if self.worker_id == 0:
            self.run_reduce()
        if dist.is_initialized():
            dist.barrier(group=comm_group)

    @staticmethod
    def output_index_to_sample_percentile(index_to_sample_fname, index_to_metric_fname, metric_name, metric_save_path,
                                          total_num_samples, sample_idx_dtype):
        to_sample = MMapIndexedDataset(index_to_sample_fname, skip_warmup=True)
        to_metric = MMapIndexedDataset(index_to_metric_fname, skip_warmup=True)
        counts = {int(to_metric[i][0]): len(to_sample[i]) for i in range(len(to_sample))}
        assert sum(counts.values()) == total_num_samples
        step = max(1, len(to_sample) // 100)
        merged = [np.concatenate(to_sample[i:min(len(to_sample), i + step)], axis=None)
                  for i in range(0, len(to_sample), step)]
        _write_rows(f"{metric_save_path}/{metric_name}_index_to_sample_percentile_merged", merged, sample_idx_dtype)
        return counts


class DistributedDataAnalyzer:
    """Same outputs, but every rank maps its contiguous slice and the values are gathered to rank 0 with collectives
    instead of through partial files (reference ``data_analyzer.py:457``)."""

    def __init__(self, dataset, num_workers=1, num_threads=1, worker_id=0, batch_size=1, metric_names=(),
                 metric_functions=(), metric_types=(), save_path="./", collate_fn=None, device="cuda",
                 comm_group=None, sample_indices=None):
        self.dataset, self.batch_size = dataset, batch_size
        self.metric_names, self.metric_functions = list(metric_names), list(metric_functions)
        self.metric_types, self.save_path, self.collate_fn = list(metric_types), save_path, collate_fn
        self.device = device if (device != "cuda" or torch.cuda.is_available()) else "cpu"
        self.comm_group, self.sample_indices = comm_group, sample_indices
        if dist.is_initialized():
            self.num_workers, self.worker_id = dist.get_world_size(comm_group), dist.get_rank(comm_group)
        else:
            self.num_workers, self.worker_id = 1, 0
        self.num_threads = 1

    def run_map_reduce(self):
        splits, _ = split_dataset(self.dataset, self.num_workers, self.worker_id, 1)
        start, end = (int(x) for x in splits[self.worker_id])
        loader = DataLoader(Subset(self.dataset, range(start, end)), batch_size=self.batch_size, shuffle=False,
                            num_workers=0, collate_fn=self.collate_fn)
        local = [[] if t == SINGLE else None for t in self.metric_types]
        for data in loader:
            for m, (fn, mtype) in enumerate(zip(self.metric_functions, self.metric_types)):
                vals = torch.as_tensor(fn(data)).to(torch.int64).reshape(-1)
                if mtype == SINGLE:
                    local[m].append(vals)
                else:
                    local[m] = vals if local[m] is None else local[m] + vals
        for m, (name, mtype) in enumerate(zip(self.metric_names, self.metric_types)):
            mdir = f"{self.save_path}/{name}"
            if mtype == SINGLE:
                mine = (torch.cat(local[m]) if local[m] else torch.zeros(0, dtype=torch.int64)).to(self.device)
                full = Dist.gather_v(mine, 0, self.comm_group, self.num_workers, self.worker_id)
                if self.worker_id == 0:
                    reduce_single_value_metric(full.cpu().numpy(), name, mdir, len(self.dataset))
            else:
                acc = local[m].to(self.device)
                if dist.is_initialized() and self.num_workers > 1:
                    dist.all_reduce(acc, group=self.comm_group)
                if self.worker_id == 0:
                    arr = acc.cpu().numpy()
                    os.makedirs(mdir, exist_ok=True)
                    _write_rows(f"{mdir}/{name}_metric_value", [arr],
                                find_fit_int_dtype(int(arr.min()), int(arr.max())))
        if dist.is_initialized():
            dist.barrier(group=self.comm_group)


class Dist:
    """Variable-length collectives used by the distributed analyzer."""

    @staticmethod
    def min_max(tensor, comm_group):
        lo, hi = tensor.min().clone(), tensor.max().clone()
        if dist.is_initialized():
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=comm_group)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=comm_group)
        return lo, hi

    @staticmethod
    def gather_v(tensor, dst, comm_group, num_workers, worker_id):
        """Concatenation (in rank order) of every rank's 1-D tensor, on ``dst``; other ranks get an empty tensor."""
        if num_workers == 1 or not dist.is_initialized():
            return tensor
        n = torch.tensor([tensor.numel()], dtype=torch.int64, device=tensor.device)
        sizes = [torch.zeros_like(n) for _ in range(num_workers)]
        dist.all_gather(sizes, n, group=comm_group)
        mx = int(max(int(s) for s in sizes))
        padded = tensor.new_zeros(mx)
        padded[:tensor.numel()] = tensor
        bufs = [torch.empty_like(padded) for _ in range(num_workers)]
        dist.all_gather(bufs, padded, group=comm_group)
        if worker_id != dst:
            return tensor.new_zeros(0)
        return torch.cat([b[:int(s)] for b, s in zip(bufs, sizes)])

    @staticmethod
    def sample_sort(tensor, comm_group, num_workers, n_samples=100):
        """Distributed sort by sampling splitters: returns this rank's sorted bucket of the global data."""
        if num_workers == 1 or not dist.is_initialized():
            return tensor.sort().values
        tensor = tensor.sort().values
        pick = torch.linspace(0, max(tensor.numel() - 1, 0), n_samples, device=tensor.device).long()
        cand = tensor[pick] if tensor.numel() else tensor.new_zeros(n_samples)
        allc = [torch.empty_like(cand) for _ in range(num_workers)]
        dist.all_gather(allc, cand, group=comm_group)
        allc = torch.cat(allc).sort().values
        splitters = allc[torch.arange(1, num_workers, device=tensor.device) * n_samples]
        bucket = torch.bucketize(tensor, splitters, right=True)
        send = [tensor[bucket == r] for r in range(num_workers)]
        counts = torch.tensor([s.numel() for s in send], dtype=torch.int64, device=tensor.device)
        recv_counts = torch.empty_like(counts)
        dist.all_to_all_single(recv_counts, counts, group=comm_group)
        out = tensor.new_empty(int(recv_counts.sum()))
        dist.all_to_all_single(out, torch.cat(send), output_split_sizes=recv_counts.tolist(),
                               input_split_sizes=counts.tolist(), group=comm_group)
        return out.sort().values
