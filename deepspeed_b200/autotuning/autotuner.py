"""``Autotuner``: searches ZeRO stage, micro-batch size and ZeRO knobs for the best throughput (reference
``autotuning/autotuner.py:42``).

Algorithm (same as the reference): (1) a *model-info* profile run gives parameter count and activation memory per
micro-batch; (2) for each ZeRO stage whose model states fit in GPU memory, the feasible micro-batch-size range is
found by doubling until failure, a few sizes in the range are timed, and (3) around the best size the stage's tuning
space is explored by the chosen tuner; stages are visited 0/1 -> 2 -> 3 and the search stops when a higher stage no
longer improves ("fast" mode). The best config is written to ``<results_dir>/ds_config_optimal.json``."""
import copy
import json
import os

from ..utils.logging import logger
from .config import DeepSpeedAutotuningConfig
from .constants import DEFAULT_TUNING_SPACES, METRIC_PERCENT_DIFF_CONST
from .scheduler import ResourceManager
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner
from .utils import canonical_name, get_all_configs, memory_to_string, number_to_string, replace_dict

ZERO_STAGES = (0, 1, 2, 3)


class Autotuner:

    def __init__(self, args, active_resources, user_config=None, run_fn=None, gpu_memory=None):
        """``args``: launcher args (``user_script``, ``user_args``, ``hostfile``, ``master_port``) or None with
        ``user_config``; ``active_resources``: {host: [gpu ids]}; ``run_fn``: see :class:`ResourceManager`."""
        self.args = args
        self.user_config = copy.deepcopy(user_config) if user_config is not None else self._get_user_config(args)
        assert self.user_config is not None, "DeepSpeed configuration is not provided"
        self.autotuning_config = DeepSpeedAutotuningConfig(**(self.user_config.get("autotuning", {}) or {}))
        self.results_dir = self.autotuning_config.results_dir
        self.exps_dir = self.autotuning_config.exps_dir
        os.makedirs(self.results_dir, exist_ok=True)
        os.makedirs(self.exps_dir, exist_ok=True)
        self.active_resources = active_resources
        hosts = list(active_resources)
        self.exp_num_nodes = len(hosts)
        self.exp_num_gpus = min(len(v) for v in active_resources.values())
        self.rm = ResourceManager(args, hosts, self.exp_num_gpus, self.results_dir, self.exps_dir,
                                  self.autotuning_config.arg_mappings, run_fn=run_fn)
        self.gpu_mem = gpu_memory if gpu_memory is not None else self.get_gpu_memory_info()
        self.records = {}
        self.optimal_cmd = None
        self.optimal_ds_config = None
        self.model_info = self.autotuning_config.model_info
        self.start_time = None

    # ---- inputs ---------------------------------------------------------------------------------
    @staticmethod
    def _get_user_config(args):
        user_args = list(getattr(args, "user_args", []) or [])
        for i, a in enumerate(user_args):
            if a in ("--deepspeed_config", "--deepspeed", "--ds_config") and i + 1 < len(user_args):
                path = user_args[i + 1]
                if path.endswith(".json") and os.path.exists(path):
                    with open(path) as f:
                        return json.load(f)
        return None

    def metric(self):
        return self.autotuning_config.metric

    def fast_enabled(self):
        return self.autotuning_config.fast

    def mp_size(self):
        return self.autotuning_config.mp_size

    def max_train_batch_size(self):
        return self.autotuning_config.max_train_batch_size

    def max_train_micro_batch_size_per_gpu(self):
        """Upper end of the micro-batch search; a user-given ``max_train_batch_size`` caps it (at least one
        gradient-accumulation step must fit)."""
        cap = self.autotuning_config.max_train_micro_batch_size_per_gpu
        mtb = self.max_train_batch_size()
        if mtb and mtb > 0:
            return min(cap, mtb * self.mp_size() // (self.exp_num_gpus * self.exp_num_nodes))
        return cap

    def min_train_micro_batch_size_per_gpu(self):
        return self.autotuning_config.min_train_micro_batch_size_per_gpu

    def num_tuning_micro_batch_sizes(self):
        return self.autotuning_config.num_tuning_micro_batch_sizes

    def get_plateau_mbs(self, tuning_space_name):
        """Largest micro-batch size of a tuning space before the metric stops improving (drops, or gains < 5 %)."""
        recs = sorted(self.records.get(tuning_space_name, []),
                      key=lambda r: r[0]["ds_config"].get("train_micro_batch_size_per_gpu", 0))
        prev_val, prev_mbs = None, 0
        for exp, val, _ in recs:
            if prev_val:
                if val < prev_val or (val - prev_val) / prev_val < 0.05:
                    break
            prev_val, prev_mbs = val, exp["ds_config"].get("train_micro_batch_size_per_gpu", 0)
        return prev_mbs

    def get_val_from_user_args(self, ds_name):
        """Numeric value the user script's command line gives to the argument mapped to ``ds_name``
        (``autotuning.arg_mappings``), or None."""
        mappings = self.autotuning_config.arg_mappings
        user_args = list(getattr(self.args, "user_args", []) or [])
        if mappings and ds_name in mappings and mappings[ds_name] in user_args:
            idx = user_args.index(mappings[ds_name])
            if idx + 1 < len(user_args) and str(user_args[idx + 1]).isnumeric():
                return user_args[idx + 1]
        return None

    def fp16_enabled(self):
        return bool(self.user_config.get("fp16", {}).get("enabled", False)) or \
            bool(self.user_config.get("bf16", {}).get("enabled", False))

    def get_gpu_memory_info(self):
        try:
            import torch
            if torch.cuda.is_available():
                return torch.cuda.get_device_properties(0).total_memory
        except Exception:  # noqa: BLE001
            pass
        return 180 * 1024**3

    def get_model_num_params(self):
        return (self.model_info or {}).get("num_params")

    def get_activation_memory_per_gpu(self):
        return (self.model_info or {}).get("activation_mem_per_gpu", 0)

    def get_instantiation_memory_required_per_gpu(self, zero_stage):
        """params + grads + optimizer state bytes per GPU for a ZeRO stage (mixed precision: 2 + 2 + 12 per param)."""
        n = self.get_model_num_params()
        gpus = self.exp_num_nodes * self.exp_num_gpus // self.mp_size()
        mixed = self.fp16_enabled()
        params = n * (2 if mixed else 4)
        grads = n * (2 if mixed else 4)
        optim = n * (12 if mixed else 8)
        if zero_stage >= 1:
            optim /= gpus
        if zero_stage >= 2:
            grads /= gpus
        if zero_stage >= 3:
            params /= gpus
        return (params + grads + optim) / self.mp_size()

    # ---- experiments ----------------------------------------------------------------------------
    def _exp(self, ds_config, name):
        return {"name": name, "ds_config": ds_config, "num_gpus": self.exp_num_gpus, "num_nodes": self.exp_num_nodes,
                "overwrite": self.autotuning_config.overwrite}

    def run_ds_config(self, ds_config, exp_name):
        exp = self._exp(copy.deepcopy(ds_config), exp_name)
        self.rm.schedule_experiments([exp])
        self.rm.run()
        val, err = None, None
        for _, (e, er) in self.rm.finished_experiments.items():
            val, err = self.rm.metric_of(e, self.metric()), er
        self.rm.clear()
        return None if err else val

    def model_info_profile_run(self):
        """ZeRO-3, micro-batch 1 run whose only job is to write ``model_info.json`` (num_params, activation memory)."""
        cfg = replace_dict(copy.deepcopy(self.user_config), {"train_micro_batch_size_per_gpu": 1,
                                                             "zero_optimization": {"stage": 3}})
        cfg.setdefault("autotuning", {})["model_info"] = {"profile": True}
        exp = self._exp(cfg, "profile_model_info")
        self.rm.schedule_experiments([exp])
        self.rm.run()
        info = None
        for _, (e, err) in self.rm.finished_experiments.items():
            path = os.path.join(e["result_dir"], "model_info.json")
            if not err and os.path.exists(path):
                with open(path) as f:
                    info = json.load(f)
            elif not err and self.rm.metric_of(e, "num_params") is not None:
                info = dict(self.rm._metrics[e["name"]])
        self.rm.clear()
        return info

    def _mbs_config(self, stage, mbs, extra=None):
        cfg = copy.deepcopy(self.user_config)
        gas = self.get_gas_from_user_config()
        replace_dict(cfg, {"train_micro_batch_size_per_gpu": mbs, "gradient_accumulation_steps": gas,
                           "zero_optimization": {"stage": stage}})
        cfg.pop("train_batch_size", None)
        if extra:
            replace_dict(cfg, extra)
        return cfg

    def get_gas_from_user_config(self):
        gas = self.user_config.get("gradient_accumulation_steps", 1)
        return 1 if gas == "auto" else int(gas)

    def get_min_max_micro_batch_size(self, stage, min_mbs, calculated_max_mbs):
        """Doubling search for the largest micro-batch size that runs (reference :851)."""
        if self.run_ds_config(self._mbs_config(stage, min_mbs), f"z{stage}_mbs{min_mbs}_probe") is None:
            return -1, -1
        lo, hi = min_mbs, max(min_mbs, calculated_max_mbs)
        # grow while it runs
        cur = lo
        while cur * 2 <= min(hi * 2, self.autotuning_config.max_train_micro_batch_size_per_gpu):
            nxt = cur * 2
            if self.run_ds_config(self._mbs_config(stage, nxt), f"z{stage}_mbs{nxt}_probe") is None:
                break
            cur = nxt
        return min_mbs, cur

    def get_tuning_micro_batch_size_list(self, min_mbs, max_mbs, num):
        if min_mbs <= 0 or max_mbs < min_mbs:
            return []
        if num <= 1 or max_mbs == min_mbs:
            return [max_mbs]
        out = {min_mbs, max_mbs}
        step = (max_mbs - min_mbs) / (num - 1)
        for i in range(1, num - 1):
            out.add(int(round(min_mbs + i * step)))
        return sorted(out)

    def run_tuning_micro_batch_sizes(self, sizes, stage):
        best_mbs, best_val = None, None
        for mbs in sizes:
            name = f"z{stage}_mbs{mbs}_gas{self.get_gas_from_user_config()}"
            val = self.run_ds_config(self._mbs_config(stage, mbs), name)
            self.update_records(f"z{stage}", {"name": name, "ds_config": self._mbs_config(stage, mbs)}, val, 1)
            if val is not None and (best_val is None or val > best_val):
                best_mbs, best_val = mbs, val
        return best_mbs, best_val

    def tune_space(self, stage, best_mbs):
        space = copy.deepcopy(DEFAULT_TUNING_SPACES[stage])
        user_space = self.user_config.get("autotuning", {}).get("tuning_space")
        if user_space:
            replace_dict(space, user_space)
        configs = get_all_configs(space)
        if len(configs) <= 1:
            return
        keys = sorted({k for c in configs for k in self._flat(c)})
        exps = []
        for c in configs[:self.autotuning_config.tuner_num_trials * 4]:
            cfg = self._mbs_config(stage, best_mbs, c)
            exps.append(self._exp(cfg, canonical_name(cfg, keys)))
        kind = self.autotuning_config.tuner_type
        cls = {"gridsearch": GridSearchTuner, "random": RandomTuner, "model_based": ModelBasedTuner}[kind]
        tuner = cls(exps, self.rm, self.metric())
        n = tuner.tune(sample_size=max(1, len(self.rm.nodes)), n_trials=self.autotuning_config.tuner_num_trials,
                       early_stopping=self.autotuning_config.tuner_early_stopping)
        if tuner.best_exp is not None:
            self.update_records(f"z{stage}", tuner.best_exp, tuner.best_metric_val, n)

    @staticmethod
    def _flat(d):
        from .utils import _flatten
        return _flatten(d)

    def update_records(self, space_name, exp, metric_val, num_exps):
        rec = self.records.setdefault(space_name, [])
        rec.append((exp, metric_val, num_exps))

    def get_best_space_record(self, space_name):
        best = None
        for exp, val, n in self.records.get(space_name, []):
            if val is not None and (best is None or val > best[1]):
                best = (exp, val, n)
        return best

    def get_best_space_records(self):
        out, overall = {}, None
        for name in self.records:
            b = self.get_best_space_record(name)
            if b is not None:
                out[name] = b
                if overall is None or b[1] > overall[1]:
                    overall = b
        if overall is not None:
            out["global"] = overall
        return out

    # ---- main -----------------------------------------------------------------------------------
    def tune(self):
        import time
        self.start_time = time.time()
        if not self.model_info or not self.model_info.get("num_params"):
            info = self.model_info_profile_run()
            if info is not None:
                self.model_info = info
        if not self.model_info or not self.model_info.get("num_params"):
            logger.warning("autotuning: model info unavailable, all ZeRO stages will be probed")
        user_stage = self.user_config.get("zero_optimization", {}).get("stage", None)
        stages = [user_stage] if isinstance(user_stage, int) else list(ZERO_STAGES)
        prev_best = None
        for stage in stages:
            if self.model_info and self.model_info.get("num_params"):
                need = self.get_instantiation_memory_required_per_gpu(stage)
                if need > self.gpu_mem:
                    logger.info(f"ZeRO stage {stage}: model states need {memory_to_string(need, 'B')} > GPU memory "
                                f"{memory_to_string(self.gpu_mem, 'B')}, skipped")
                    continue
                act = self.get_activation_memory_per_gpu()
                calc_max = int((self.gpu_mem - need) // act) if act else self.autotuning_config.min_train_micro_batch_size_per_gpu
            else:
                calc_max = self.autotuning_config.min_train_micro_batch_size_per_gpu
            lo, hi = self.get_min_max_micro_batch_size(stage, self.autotuning_config.min_train_micro_batch_size_per_gpu,
                                                       max(1, calc_max))
            if lo < 0:
                continue
            sizes = self.get_tuning_micro_batch_size_list(lo, hi, self.autotuning_config.num_tuning_micro_batch_sizes)
            best_mbs, best_val = self.run_tuning_micro_batch_sizes(sizes, stage)
            if best_mbs is None:
                continue
            self.tune_space(stage, best_mbs)
            best = self.get_best_space_record(f"z{stage}")
            if self.fast_enabled() and prev_best is not None and best is not None and \
                    best[1] < prev_best * (1 - METRIC_PERCENT_DIFF_CONST) and stage >= 2:
                logger.info(f"ZeRO stage {stage} is slower than the previous stage: search stopped (fast mode)")
                break
            if best is not None:
                prev_best = best[1] if prev_best is None else max(prev_best, best[1])
        self.write_optimal_config()
        return self.get_best_space_records()

    def print_tuning_results(self):
        best = self.get_best_space_records()
        rows = [(k, n, v, e["name"]) for k, (e, v, n) in best.items()]
        try:
            from tabulate import tabulate
            print(tabulate(rows, headers=["tuning_space", "num_experiments", "best_metric_val", "best_exp_name"],
                           tablefmt="pipe"))
        except ImportError:
            for r in rows:
                print(r)
        if self.model_info and self.model_info.get("num_params"):
            print(f"model parameters: {number_to_string(self.model_info['num_params'])}")

    def write_optimal_config(self):
        best = self.get_best_space_records().get("global")
        if best is None:
            return None
        exp = best[0]
        cfg = copy.deepcopy(exp["ds_config"])
        cfg.pop("autotuning", None)
        self.optimal_ds_config = cfg
        path = os.path.join(self.results_dir, "ds_config_optimal.json")
        with open(path, "w") as f:
            json.dump(cfg, f, indent=2)
        if self.args is not None and getattr(self.args, "user_script", None):
            user_args = list(self.args.user_args)
            for i, a in enumerate(user_args):
                if a in ("--deepspeed_config", "--deepspeed", "--ds_config") and i + 1 < len(user_args):
                    user_args[i + 1] = path
            self.optimal_cmd = ["deepspeed", self.args.user_script] + user_args
            with open(os.path.join(self.results_dir, "cmd_optimal.txt"), "w") as f:
                f.write(" ".join(self.optimal_cmd))
        return path

    def run_after_tuning(self):
        if self.optimal_cmd:
            import subprocess
            import sys
            cmd = [sys.executable, "-m", "deepspeed_b200.launcher.runner"] + self.optimal_cmd[1:]
            return subprocess.Popen(cmd).wait()
        logger.info("No optimal DeepSpeed configuration found by autotuning.")
        return None
