"""Experiment scheduler (reference ``autotuning/scheduler.py:32`` ``ResourceManager``): runs each experiment - a
ds_config variant of the user's training command - on a free slice of the reserved nodes/GPUs, collects the metric
file the engine writes at ``end_profile_step`` (or the error from stderr.log).

``run_fn`` (optional) replaces the subprocess launch by a Python callable ``run_fn(exp) -> dict of metrics`` - used by
tests and by in-process tuning of short jobs."""
import copy
import json
import os
import subprocess
import sys
import time

from ..utils.logging import logger
from .utils import search_error

TIMEOUT = 5


class Node:

    def __init__(self, host, max_slots):
        self.host, self.max_slots = host, max_slots
        self.idle_slots = list(range(max_slots))

    def reserve_slots(self, slot_request):
        if len(self.idle_slots) >= slot_request:
            return [self.idle_slots.pop(0) for _ in range(slot_request)]
        return None

    def restore_slots(self, slots):
        self.idle_slots += slots


class Reservation:

    def __init__(self, node, slots):
        self.node, self.slots = node, slots

    def restore_slots(self):
        self.node.restore_slots(self.slots)

    def desc(self):
        return f"{self.node.host}:{','.join(map(str, self.slots))}@"


class ResourceManager:

    def __init__(self, args, hosts, num_gpus_per_node, results_dir, exps_dir, arg_mappings=None, run_fn=None):
        self.results_dir, self.exps_dir = results_dir, exps_dir
        self.nodes = [Node(h, num_gpus_per_node) for h in hosts]
        self.experiment_queue = []
        self.running_experiments = {}
        self.finished_experiments = {}
        self.experiment_count = 0
        self.exp_paths = set()
        self.args = args
        self.arg_mappings = {k.strip(): v.strip() for k, v in (arg_mappings or {}).items()}
        self.run_fn = run_fn
        self._metrics = {}

    # ---- queueing -------------------------------------------------------------------------------
    def schedule_experiments(self, exps):
        for exp in exps:
            if isinstance(exp, str):
                with open(exp) as f:
                    exp = json.load(f)
            exp = copy.deepcopy(exp)
            exp["exp_id"] = self.experiment_count
            self.experiment_count += 1
            result_dir = exp["result_dir"] = os.path.join(self.results_dir, exp["name"])
            if self.args is not None and getattr(self.args, "user_script", None):
                exp["user_script"], exp["user_args"] = self.args.user_script, list(self.args.user_args)
            exp["hostfile"] = getattr(self.args, "hostfile", None)
            exp["master_port"] = getattr(self.args, "master_port", 29500) if self.args is not None else 29500
            done = os.path.join(result_dir, "metrics.json")
            if os.path.exists(done) and not exp.get("overwrite", True):
                with open(done) as f:
                    self._metrics[exp["name"]] = json.load(f)
                self.finished_experiments[exp["exp_id"]] = (exp, None)
                continue
            self.experiment_queue.append(exp)

    def resource_request(self, exp):
        return exp.get("num_nodes", 1), exp.get("num_gpus", 1)

    def request_resources(self, exp):
        num_nodes, slots = self.resource_request(exp)
        got = []
        for node in self.nodes:
            if num_nodes == 0:
                break
            s = node.reserve_slots(slots)
            if s is not None:
                got.append(Reservation(node, s))
                num_nodes -= 1
        if num_nodes == 0:
            return got
        for r in got:
            r.restore_slots()
        return None

    # ---- running --------------------------------------------------------------------------------
    def run(self):
        while self.experiment_queue:
            exp = self.experiment_queue.pop(0)
            reservations = self.request_resources(exp)
            if reservations is None:
                self.experiment_queue.insert(0, exp)
                self._reap(block=True)
                continue
            self._launch(exp, reservations)
            self._reap(block=False)
        while self.running_experiments:
            self._reap(block=True)

    def _launch(self, exp, reservations):
        os.makedirs(exp["result_dir"], exist_ok=True)
        exp["ds_config"].setdefault("autotuning", {})
        exp["ds_config"]["autotuning"].update(enabled=True, metric_path=os.path.join(exp["result_dir"], "metrics.json"),
                                              model_info_path=os.path.join(exp["result_dir"], "model_info.json"))
        cfg_path = exp["ds_config_path"] = os.path.join(exp["result_dir"], "ds_config.json")
        with open(cfg_path, "w") as f:
            json.dump(exp["ds_config"], f)
        with open(os.path.join(exp["result_dir"], "exp.json"), "w") as f:
            json.dump({k: v for k, v in exp.items() if k != "proc"}, f, default=str)
        if self.run_fn is not None:
            err = None
            try:
                metrics = self.run_fn(exp)
                if metrics is None:
                    err = "experiment failed"
                else:
                    self._metrics[exp["name"]] = metrics
                    with open(exp["ds_config"]["autotuning"]["metric_path"], "w") as f:
                        json.dump(metrics, f)
            except Exception as e:  # noqa: BLE001
                err = f"{type(e).__name__}: {e}"
            for r in reservations:
                r.restore_slots()
            self.finished_experiments[exp["exp_id"]] = (exp, err)
            return
        include = "".join(r.desc() for r in reservations)[:-1]
        user_args = self._substitute(list(exp.get("user_args", [])), exp)
        cmd = [sys.executable, "-m", "deepspeed_b200.launcher.runner", "--master_port", str(exp["master_port"])]
        if exp.get("hostfile") and os.path.exists(exp["hostfile"]):
            cmd += ["--hostfile", exp["hostfile"], "--include", include]
        else:
            cmd += ["--num_gpus", str(len(reservations[0].slots))]
        cmd += [exp["user_script"]] + user_args
        out = open(os.path.join(exp["result_dir"], "stdout.log"), "w")
        err = open(os.path.join(exp["result_dir"], "stderr.log"), "w")
        proc = subprocess.Popen(cmd, stdout=out, stderr=err)
        self.running_experiments[exp["exp_id"]] = (exp, reservations, proc, time.time())

    def _substitute(self, user_args, exp):
        """Point the user's ``--deepspeed_config``-style argument at this experiment's config and apply
        ``arg_mappings`` (ds_config key -> command-line flag)."""
        for i, a in enumerate(user_args):
            if a in ("--deepspeed_config", "--deepspeed", "--ds_config") and i + 1 < len(user_args) \
                    and user_args[i + 1].endswith(".json"):
                user_args[i + 1] = exp["ds_config_path"]
        from .utils import _flatten
        flat = _flatten(exp["ds_config"])
        for key, flag in self.arg_mappings.items():
            if key in flat and flag in user_args:
                user_args[user_args.index(flag) + 1] = str(flat[key])
        return user_args

    def _reap(self, block):
        while True:
            for exp_id, (exp, reservations, proc, t0) in list(self.running_experiments.items()):
                if proc.poll() is not None:
                    for r in reservations:
                        r.restore_slots()
                    err = search_error(os.path.join(exp["result_dir"], "stderr.log")) if proc.returncode else None
                    mp = exp["ds_config"]["autotuning"]["metric_path"]
                    if os.path.exists(mp):
                        with open(mp) as f:
                            self._metrics[exp["name"]] = json.load(f)
                    elif err is None:
                        err = "no metric file written"
                    self.finished_experiments[exp_id] = (exp, err)
                    del self.running_experiments[exp_id]
                    return
            if not block or not self.running_experiments:
                return
            time.sleep(TIMEOUT if self.run_fn is None else 0)

    # ---- results --------------------------------------------------------------------------------
    def metric_of(self, exp, metric):
        m = self._metrics.get(exp["name"])
        return None if m is None else m.get(metric)

    def parse_results(self, metric):
        """Best finished experiment by ``metric`` (higher is better; latency is negated)."""
        best_exp, best_val = None, None
        for exp_id, (exp, err) in self.finished_experiments.items():
            if err:
                logger.info(f"The experiment exp_id = {exp_id}, exp_name = {exp['name']}, did not run successfully: {err}")
                continue
            val = self.metric_of(exp, metric)
            if val is None:
                continue
            val = -val if metric == "latency" else val
            if best_val is None or val > best_val:
                best_exp, best_val = exp, val
        return best_exp, best_val

    def clear(self):
        self.experiment_queue = []
        for exp_id, (exp, reservations, proc, _) in list(self.running_experiments.items()):
            proc.terminate()
            for r in reservations:
                r.restore_slots()
        self.running_experiments = {}
        self.finished_experiments = {}
        self.exp_paths = set()
