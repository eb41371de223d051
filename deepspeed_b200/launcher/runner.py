"""``deepspeed`` CLI front-end: resource discovery, include/exclude filtering, single- and multi-node command lines.

Reference: ``launcher/runner.py`` (``parse_args`` :48, ``fetch_hostfile`` :230, ``parse_inclusion_exclusion`` :391,
``encode_world_info`` :401, ``main`` :436-642) and ``launcher/multinode_runner.py`` (PDSH/OpenMPI/MPICH/IMPI/
MVAPICH/Slurm command builders).
"""
import argparse
import base64
import collections
import json
import os
import re
import shlex
import signal
import subprocess
import sys
from copy import deepcopy

from ..utils.logging import logger

DLTS_HOSTFILE = "/job/hostfile"
EXPORT_ENVS = ["MLFLOW", "NCCL", "PYTHON", "MV2", "UCX", "CUDA", "TORCH", "DSB200", "DS_"]
DEEPSPEED_ENVIRONMENT_NAME = os.getenv("DS_ENV_FILE", ".deepspeed_env")
PDSH_MAX_FAN_OUT = 1024
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500


def parse_args(args=None):
    p = argparse.ArgumentParser(description="DeepSpeed-B200 runner: launch multi-GPU / multi-node training jobs.")
    p.add_argument("-H", "--hostfile", type=str, default=DLTS_HOSTFILE, help="Hostfile path: lines 'host slots=N'")
    p.add_argument("-i", "--include", type=str, default="", help="e.g. 'worker-0@worker-1:0,2'")
    p.add_argument("-e", "--exclude", type=str, default="")
    p.add_argument("--num_nodes", type=int, default=-1)
    p.add_argument("--min_elastic_nodes", type=int, default=-1)
    p.add_argument("--max_elastic_nodes", type=int, default=-1)
    p.add_argument("--num_gpus", "--num_accelerators", type=int, default=-1)
    p.add_argument("--master_port", default=TORCH_DISTRIBUTED_DEFAULT_PORT, type=int)
    p.add_argument("--master_addr", default="", type=str)
    p.add_argument("--node_rank", default=-1, type=int)
    p.add_argument("--launcher", default="pdsh", type=str,
                   help="pdsh | openmpi | mpich | impi | mvapich | slurm")
    p.add_argument("--launcher_args", default="", type=str)
    p.add_argument("--module", action="store_true")
    p.add_argument("--no_python", action="store_true")
    p.add_argument("--no_local_rank", action="store_true")
    p.add_argument("--no_ssh", action="store_true")
    p.add_argument("--no_ssh_check", action="store_true")
    p.add_argument("--force_multi", action="store_true")
    p.add_argument("--save_pid", action="store_true")
    p.add_argument("--enable_each_rank_log", default="None", type=str)
    p.add_argument("--autotuning", default="", choices=["tune", "run", ""], type=str)
    p.add_argument("--elastic_training", action="store_true")
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--bind_core_list", type=str, default=None)
    p.add_argument("--ssh_port", type=int, default=None)
    p.add_argument("--venv_script", type=str, default=None)
    p.add_argument("user_script", type=str)
    p.add_argument("user_args", nargs=argparse.REMAINDER)
    return p.parse_args(args=args)


def fetch_hostfile(hostfile_path):
    if not os.path.isfile(hostfile_path):
        logger.warning("Unable to find hostfile, will proceed with training with local resources only.")
        return None
    with open(hostfile_path) as fd:
        return _parse_hostfile(fd.readlines())


def _parse_hostfile(lines):
    pattern = r"^(\S+)\s+slots=(\d+)"
    pool = collections.OrderedDict()
    for line in lines:
        line = line.strip()
        if not line or line.startswith("#"):
            continue
        m = re.search(pattern, line)
        if not m:
            raise ValueError(f"Hostfile contains a bad entry: {line}, unable to proceed with training.")
        host, slots = m.group(1), int(m.group(2))
        if host in pool:
            raise ValueError(f"Hostfile contains multiple entries for {host}, unable to proceed with training.")
        pool[host] = slots
    if not pool:
        raise ValueError("Hostfile is empty or not formatted correctly, unable to proceed with training.")
    return pool


def _stable_remove_duplicates(data):
    out = []
    for x in data:
        if x not in out:
            out.append(x)
    return out


def parse_resource_filter(host_info, include_str="", exclude_str=""):
    """``host_info``: {host: [slot ids]}. Filter strings: ``host1@host2:0,2`` (NODE_SEP '@', SLOT_LIST_START ':')."""
    if include_str and exclude_str:
        raise ValueError("include_str and exclude_str are mutually exclusive.")
    if not include_str and not exclude_str:
        return host_info
    filtered = dict() if include_str else deepcopy(host_info)
    parse_str = include_str or exclude_str
    for node_config in parse_str.split("@"):
        if ":" in node_config:
            hostname, slots = node_config.split(":")
            slots = [int(x) for x in slots.split(",")]
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            for s in slots:
                if s not in host_info[hostname]:
                    raise ValueError(f"No slot '{s}' specified on host '{hostname}'")
            if include_str:
                filtered.setdefault(hostname, [])
                filtered[hostname].extend(slots)
            else:
                for s in slots:
                    filtered[hostname].remove(s)
        else:
            hostname = node_config
            if hostname not in host_info:
                raise ValueError(f"Hostname '{hostname}' not found in hostfile")
            if include_str:
                filtered[hostname] = host_info[hostname]
            else:
                filtered[hostname] = []
    ordered = collections.OrderedDict()
    for host in host_info:
        if host in filtered and filtered[host]:
            ordered[host] = sorted(_stable_remove_duplicates(filtered[host]))
    return ordered


def parse_inclusion_exclusion(resource_pool, inclusion, exclusion):
    active = collections.OrderedDict((h, list(range(s))) for h, s in resource_pool.items())
    return parse_resource_filter(active, include_str=inclusion, exclude_str=exclusion)


def encode_world_info(world_info):
    return base64.urlsafe_b64encode(json.dumps(world_info).encode("utf-8")).decode("utf-8")


def decode_world_info(encoded):
    return json.loads(base64.urlsafe_b64decode(encoded))


def _local_gpu_count():
    try:
        import torch
        n = torch.cuda.device_count()
        if n:
            return n
    except Exception:
        pass
    return max(1, int(os.environ.get("DS_LOCAL_SLOTS", "1")))


def build_launch_command(args, active_resources, master_addr=None):
    """Per-node command: ``python -u -m deepspeed_b200.launcher.launch --world_info=... user_script args``."""
    world_info = encode_world_info(active_resources)
    cmd = [sys.executable, "-u", "-m", "deepspeed_b200.launcher.launch", f"--world_info={world_info}",
           f"--master_addr={master_addr or args.master_addr or '127.0.0.1'}", f"--master_port={args.master_port}"]
    if args.node_rank >= 0:
        cmd.append(f"--node_rank={args.node_rank}")
    for flag in ("no_python", "module", "no_local_rank", "save_pid", "elastic_training", "bind_cores_to_rank"):
        if getattr(args, flag):
            cmd.append(f"--{flag}")
    if args.enable_each_rank_log != "None":
        cmd.append(f"--enable_each_rank_log={args.enable_each_rank_log}")
    if args.bind_core_list:
        cmd.append(f"--bind_core_list={args.bind_core_list}")
    return cmd + [args.user_script] + list(args.user_args)


def _collect_exports():
    exports = {}
    for var, val in os.environ.items():
        if any(var.startswith(name) for name in EXPORT_ENVS):
            exports[var] = val
    for env_dir in (os.path.expanduser("~"), "."):
        path = os.path.join(env_dir, DEEPSPEED_ENVIRONMENT_NAME)
        if os.path.isfile(path):
            with open(path) as fd:
                for line in fd:
                    if "=" in line:
                        k, v = line.strip().split("=", 1)
                        exports[k] = v
    return exports


def parse_node_config(node_config):
    """``"worker-1:0,2"`` -> ``("worker-1", [0, 2])``; a bare host name means all of its slots (``[]``)."""
    if ":" not in node_config:
        return node_config, []
    host, slots = node_config.split(":")
    return host, [int(x) for x in slots.split(",")]


def parse_node_config_list(node_config_list):
    """``"worker-0@worker-1:0,2"`` -> ``{host: sorted slot list}`` (the include / exclude string syntax)."""
    out = collections.defaultdict(list)
    for item in node_config_list.split("@"):
        host, slots = parse_node_config(item)
        out[host] += slots
    return {h: sorted(set(v)) for h, v in out.items()}


def parse_num_nodes(str_num_nodes, elastic_training):
    """``"4"`` -> (4, -1); ``"2:8"`` (MIN:MAX) only with elastic training."""
    parts = str(str_num_nodes).split(":")
    if len(parts) == 1:
        return int(parts[0]), -1
    if len(parts) == 2 and elastic_training:
        return int(parts[0]), int(parts[1])
    if len(parts) == 2:
        raise RuntimeError("MIN:MAX format is only supported in elastic training")
    raise RuntimeError(f"num_nodes {str_num_nodes} is not in MIN:MAX format")


def run_autotuning(args, active_resources):
    """``deepspeed --autotuning tune|run script.py --deepspeed_config ds.json`` (reference ``runner.py:407``)."""
    from ..autotuning import Autotuner
    tuner = Autotuner(args, active_resources)
    logger.info("[Start] Running autotuning")
    tuner.tune()
    tuner.print_tuning_results()
    logger.info("[End] Running autotuning")
    tuner.write_optimal_config()
    if args.autotuning == "run":
        tuner.run_after_tuning()


def main(args=None):
    args = parse_args(args)
    resource_pool = fetch_hostfile(args.hostfile)
    if not resource_pool:
        n = args.num_gpus if args.num_gpus > 0 else _local_gpu_count()
        resource_pool = collections.OrderedDict(localhost=n)
        args.master_addr = args.master_addr or "127.0.0.1"
    active = parse_inclusion_exclusion(resource_pool, args.include, args.exclude)
    if args.num_nodes > 0:
        active = collections.OrderedDict(list(active.items())[:args.num_nodes])
    if args.num_gpus > 0:
        active = collections.OrderedDict((h, list(range(args.num_gpus))) for h in active)
    if args.autotuning:
        run_autotuning(args, active)
        return
    multi_node = args.force_multi or len(active) > 1
    if not args.master_addr:
        first = list(active.keys())[0]
        args.master_addr = "127.0.0.1" if first == "localhost" else first
    env = os.environ.copy()
    if not multi_node or args.no_ssh:
        cmd = build_launch_command(args, active)
    else:
        from .multinode_runner import get_runner
        runner = get_runner(args.launcher, args, encode_world_info(active), active)
        if not runner.backend_exists():
            raise RuntimeError(f"launcher '{args.launcher}' not installed.")
        for k, v in _collect_exports().items():
            runner.add_export(k, v)
        cmd, kill_cmd, env = runner.get_cmd(env, active)
    logger.info(f"cmd = {' '.join(shlex.quote(c) for c in cmd)}")
    result = subprocess.Popen(cmd, env=env)

    def sigkill_handler(signum, frame):
        result.send_signal(signal.SIGINT)
        import time
        time.sleep(0.1)
        result.send_signal(signal.SIGTERM)
        result.wait()
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    result.wait()
    if result.returncode > 0:
        sys.exit(result.returncode)


if __name__ == "__main__":
    main()
