"""Curriculum-learning batch sampler (reference ``runtime/data_pipeline/data_sampling/data_sampler.py:36``).

Every curriculum step each metric's scheduler yields a difficulty; when any difficulty changes, rank 0 materialises the
newly admitted samples as a shuffled *cluster* file (mmap indexed dataset), then draws each global batch from all
clusters proportionally to their size and broadcasts the indices; every DP rank takes its contiguous slice of each
micro-batch. State (clusters on disk + positions + rng) is checkpointable."""
import os

import numpy as np
import torch

from .... import comm as dist
from ....utils.logging import logger
from ..curriculum_scheduler import CurriculumScheduler
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype


def _device():
    return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")


class DeepSpeedDataSampler:

    def __init__(self, data_efficiency_config, one_epoch_total_samples, micro_batch_size, data_parallel_rank,
                 data_parallel_size, data_parallel_group, gradient_accumulation_steps, global_rank, drop_last=True):
        self.data_efficiency_config = data_efficiency_config
        ds_cfg = data_efficiency_config["data_sampling"]
        self.cl_cfg = ds_cfg.get("curriculum_learning", {"enabled": False})
        self.one_epoch_total_samples = one_epoch_total_samples
        self.index_dtype = find_fit_int_dtype(0, one_epoch_total_samples)
        self.total_samples = one_epoch_total_samples * ds_cfg.get("num_epochs", 1000)
        self.micro_batch_size = micro_batch_size
        self.data_parallel_rank = data_parallel_rank
        self.data_parallel_size = data_parallel_size
        self.data_parallel_group = data_parallel_group
        self.micro_batch_times_data_parallel_size = micro_batch_size * data_parallel_size
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.global_batch_size = self.micro_batch_times_data_parallel_size * gradient_accumulation_steps
        self.global_rank = global_rank
        self.drop_last = drop_last
        self.np_rng = np.random.default_rng(data_efficiency_config.get("seed", 1234))
        self.batch = []
        self.consumed_samples = 0
        self.curriculum_enabled = bool(self.cl_cfg.get("enabled", False))
        self.curriculum_step = 0
        self.current_difficulties = {}
        self.data_cluster_paths = []
        self.data_cluster_current_position = []
        self.data_clusters, self.data_cluster_sizes = [], []
        self.curriculum_schedulers, self.difficulty_type, self.clustering_type = {}, {}, {}
        self.curriculum_index_to_sample, self.curriculum_index_to_metric = {}, {}
        self.data_1epoch_size = None
        if self.curriculum_enabled:
            if global_rank == 0:
                os.makedirs(self.cl_cfg["data_cluster_path"], exist_ok=True)
            for metric, mcfg in self.cl_cfg["curriculum_metrics"].items():
                self.curriculum_schedulers[metric] = CurriculumScheduler(mcfg)
                self.difficulty_type[metric] = mcfg["difficulty_type"]
                self.clustering_type[metric] = mcfg["clustering_type"]
                if global_rank == 0 and mcfg["clustering_type"] != "single_cluster":
                    self.curriculum_index_to_sample[metric] = MMapIndexedDataset(mcfg["index_to_sample_path"],
                                                                                 skip_warmup=True)
                    if mcfg["difficulty_type"] == "value":
                        self.curriculum_index_to_metric[metric] = MMapIndexedDataset(mcfg["index_to_metric_path"],
                                                                                     skip_warmup=True)
        assert self.total_samples > 0 and micro_batch_size > 0 and data_parallel_size > 0
        assert data_parallel_rank < data_parallel_size

    def __len__(self):
        return self.total_samples

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        for metric, sched in self.curriculum_schedulers.items():
            if metric in schedule_func_dict:
                sched.set_custom_get_difficulty(schedule_func_dict[metric])

    def get_start_end_idx(self, batch_len=None):
        batch_len = batch_len or self.micro_batch_times_data_parallel_size
        at = lambda r: round(r * batch_len / self.data_parallel_size)  # noqa: E731
        return at(self.data_parallel_rank), at(self.data_parallel_rank + 1)

    # ---- which samples does a difficulty interval admit -------------------------------------
    def get_sample_based_on_metric_value(self, metric, value_start, value_end):
        to_sample, to_metric = self.curriculum_index_to_sample[metric], self.curriculum_index_to_metric[metric]
        picked = [np.copy(to_sample[row]) for row in range(len(to_sample))
                  if value_start < to_metric[row] <= value_end]
        return np.concatenate(picked, axis=None) if picked else None

    def get_sample_based_on_metric_percentile(self, metric, percentile_start, percentile_end):
        to_sample = self.curriculum_index_to_sample[metric]
        if self.data_1epoch_size is None:
            self.data_1epoch_size = int(sum(len(x) for x in to_sample))
        max_percentile = self.cl_cfg["curriculum_metrics"][metric]["max_difficulty"]
        per = self.data_1epoch_size // max_percentile
        start_count, end_count = per * percentile_start, per * percentile_end
        if percentile_end == max_percentile:
            end_count = self.data_1epoch_size
        picked, seen = [], 0
        for row in range(len(to_sample)):
            n = len(to_sample[row])
            if seen + n > start_count:
                lo = max(0, start_count - seen)
                hi = n if seen + n <= end_count else end_count - seen
                picked.append(np.copy(to_sample[row][lo:hi]))
            seen += n
            if seen >= end_count:
                break
        return np.concatenate(picked, axis=None) if picked else None

    def _admitted(self, metric, lo):
        hi = self.current_difficulties[metric]
        if self.difficulty_type[metric] == "value":
            return self.get_sample_based_on_metric_value(metric, lo, hi)
        return self.get_sample_based_on_metric_percentile(metric, lo, hi)

    def _cluster_file(self, fname):
        return f"{self.cl_cfg['data_cluster_path']}/{fname}"

    def get_new_cluster(self, previous_difficulties):
        fname = "cluster" + "".join(f"_{m}{self.current_difficulties[m]}" for m in self.curriculum_schedulers)
        path = self._cluster_file(fname)
        if self.global_rank == 0:
            clustered = [m for m, t in self.clustering_type.items() if t != "single_cluster"]
            new_cluster = None
            if len(clustered) > 1:
                # several metrics: intersect everything admitted so far, minus what earlier clusters already hold
                for metric in self.curriculum_schedulers:
                    if self.clustering_type[metric] == "single_cluster":
                        mc = np.arange(self.one_epoch_total_samples, dtype=self.index_dtype)
                    else:
                        mc = self._admitted(metric, float("-inf") if self.difficulty_type[metric] == "value" else 0)
                    new_cluster = mc if new_cluster is None else np.intersect1d(new_cluster, mc, assume_unique=True)
                for cluster in self.data_clusters:
                    new_cluster = np.setdiff1d(new_cluster, cluster[0], assume_unique=True)
            else:
                if not self.data_clusters:
                    new_cluster = np.arange(self.one_epoch_total_samples, dtype=self.index_dtype)
                for metric in clustered:
                    new_cluster = self._admitted(metric, previous_difficulties[metric])
            if new_cluster is not None and len(new_cluster) > 0:
                logger.info(f"new data cluster ({previous_difficulties} -> {self.current_difficulties}) "
                            f"with {len(new_cluster)} samples")
                self._write_cluster(path, new_cluster)
                self.data_clusters.append(MMapIndexedDataset(path, skip_warmup=True))
                self.data_cluster_sizes.append(len(self.data_clusters[-1][0]))
            else:
                logger.info(f"new data cluster ({previous_difficulties} -> {self.current_difficulties}) is empty")
        if dist.is_initialized():
            dist.barrier(group=self.data_parallel_group)
        if os.path.isfile(f"{path}.bin"):
            self.data_cluster_paths.append(fname)
            self.data_cluster_current_position.append(0)

    def _write_cluster(self, path, samples):
        samples = np.array(samples, dtype=self.index_dtype)
        self.np_rng.shuffle(samples)
        builder = create_mmap_dataset_builder(path, self.index_dtype)
        builder.add_item_numpy(samples)
        close_mmap_dataset_builder(builder, path)

    def sample_from_clusters(self):
        sizes = np.asarray(self.data_cluster_sizes, dtype=np.float64)
        draws = self.np_rng.choice(len(sizes), self.global_batch_size, replace=True, p=sizes / sizes.sum())
        return np.bincount(draws, minlength=len(sizes))

    def reshuffle_clusters(self, cidx):
        path = self._cluster_file(self.data_cluster_paths[cidx])
        self._write_cluster(path, np.copy(self.data_clusters[cidx][0]))
        self.data_clusters[cidx] = MMapIndexedDataset(path, skip_warmup=True)

    def get_sample_from_cluster(self, cidx, num_samples):
        start = self.data_cluster_current_position[cidx]
        samples = list(np.copy(self.data_clusters[cidx][0][start:start + num_samples]))
        self.data_cluster_current_position[cidx] += num_samples
        if len(samples) < num_samples:
            remain = num_samples - len(samples)
            self.reshuffle_clusters(cidx)
            samples += list(np.copy(self.data_clusters[cidx][0][:remain]))
            self.data_cluster_current_position[cidx] = remain
        return samples

    def get_next_global_batch(self):
        if not self.curriculum_enabled:
            # plain sequential sampling when curriculum learning is off
            start = self.consumed_samples % self.one_epoch_total_samples
            self.batch = [(start + i) % self.one_epoch_total_samples for i in range(self.global_batch_size)]
            return
        self.curriculum_step += 1
        changed, previous = False, {}
        for metric, sched in self.curriculum_schedulers.items():
            nxt = sched.update_difficulty(self.curriculum_step)
            if self.current_difficulties.get(metric) != nxt:
                changed = True
            previous[metric] = self.current_difficulties.get(
                metric, float("-inf") if self.difficulty_type[metric] == "value" else 0)
            self.current_difficulties[metric] = nxt
        if changed:
            self.get_new_cluster(previous)
        dev = _device()
        if self.global_rank == 0:
            batch = []
            for cidx, n in enumerate(self.sample_from_clusters()):
                batch += self.get_sample_from_cluster(cidx, int(n))
            self.np_rng.shuffle(batch)
            assert len(batch) <= self.global_batch_size
            batch += [-1] * (self.global_batch_size - len(batch))
            batch = torch.tensor([int(b) for b in batch], device=dev, dtype=torch.long)
        else:
            batch = torch.empty(self.global_batch_size, device=dev, dtype=torch.long)
        if dist.is_initialized() and self.data_parallel_size > 1:
            dist.broadcast(batch, dist.get_global_rank(self.data_parallel_group, 0)
                           if self.data_parallel_group is not None else 0, group=self.data_parallel_group)
        self.batch = batch[batch != -1].tolist()

    def __iter__(self):
        while self.consumed_samples <= self.total_samples:
            if len(self.batch) == 0:
                self.get_next_global_batch()
            n = self.micro_batch_times_data_parallel_size
            current, self.batch = self.batch[:n], self.batch[n:]
            if len(current) == n or (len(current) > 0 and not self.drop_last):
                s, e = self.get_start_end_idx(len(current))
                yield current[s:e]
                self.consumed_samples += len(current)

    def state_dict(self):
        return {
            "batch": list(self.batch),
            "consumed_samples": self.consumed_samples,
            "curriculum_step": self.curriculum_step,
            "current_difficulties": dict(self.current_difficulties),
            "data_cluster_paths": list(self.data_cluster_paths),
            "data_cluster_current_position": list(self.data_cluster_current_position),
            "np_rng_state": self.np_rng.bit_generator.state,
        }

    def load_state_dict(self, state_dict):
        self.batch = state_dict["batch"]
        self.consumed_samples = state_dict["consumed_samples"]
        self.curriculum_step = state_dict["curriculum_step"]
        self.current_difficulties = state_dict["current_difficulties"]
        self.data_cluster_paths = [p.split("/")[-1] for p in state_dict["data_cluster_paths"]]
        self.data_cluster_current_position = state_dict["data_cluster_current_position"]
        rng_state = state_dict["np_rng_state"]
        if isinstance(rng_state, dict):
            self.np_rng.bit_generator.state = rng_state
        else:  # reference checkpoints hold the legacy global numpy state
            np.random.set_state(rng_state)
        for metric, sched in self.curriculum_schedulers.items():
            if metric in self.current_difficulties:
                sched.set_current_difficulty(self.current_difficulties[metric])
        if self.global_rank == 0:
            self.data_clusters, self.data_cluster_sizes = [], []
            for fname in self.data_cluster_paths:
                self.data_clusters.append(MMapIndexedDataset(self._cluster_file(fname), skip_warmup=True))
                self.data_cluster_sizes.append(len(self.data_clusters[-1][0]))
