"""Per-node launcher: spawns one process per local GPU with RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* set, watches them,
kills the whole tree on the first failure and propagates its exit code.

Reference: ``launcher/launch.py`` (``main`` :145, env setup :187-192, Popen per rank :237-272, watchdog :330-372,
``terminate_process_tree`` :130, per-rank log files :225-235).
"""
import argparse
import os
import signal
import subprocess
import sys
import time

import psutil

from ..utils.logging import logger
from .runner import decode_world_info

PID_FILE_BASEPATH = "/tmp"


def parse_args(args=None):
    p = argparse.ArgumentParser()
    p.add_argument("--node_rank", type=int, default=0)
    p.add_argument("--master_addr", default="127.0.0.1", type=str)
    p.add_argument("--master_port", default=29500, type=int)
    p.add_argument("--world_info", default="None", type=str)
    p.add_argument("--module", action="store_true")
    p.add_argument("--no_python", action="store_true")
    p.add_argument("--enable_each_rank_log", default="None", type=str)
    p.add_argument("--save_pid", type=int, default=0, nargs="?", const=1)
    p.add_argument("--no_local_rank", action="store_true")
    p.add_argument("--elastic_training", action="store_true")
    p.add_argument("--bind_cores_to_rank", action="store_true")
    p.add_argument("--bind_core_list", type=str, default=None)
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return p.parse_args(args)


def terminate_process_tree(pid):
    try:
        process = psutil.Process(pid)
    except psutil.NoSuchProcess:
        return
    children = process.children(recursive=True)
    children.append(process)
    for child in children:
        try:
            child.terminate()
        except psutil.NoSuchProcess:
            pass
    gone, alive = psutil.wait_procs(children, timeout=30)
    for p in alive:
        try:
            p.kill()
        except psutil.NoSuchProcess:
            pass


def build_rank_env(world_info, node_rank, master_addr, master_port):
    """Returns (list of per-local-rank env dicts, global world size)."""
    node_list = list(world_info.keys())
    local_node = node_list[node_rank]
    local_gpu_ids = world_info[local_node]
    world_size = sum(len(v) for v in world_info.values())
    global_rank_mapping, curr = {}, 0
    for node in node_list:
        global_rank_mapping[node] = []
        for _ in world_info[node]:
            global_rank_mapping[node].append(curr)
            curr += 1
    envs = []
    for local_rank, gpu in enumerate(local_gpu_ids):
        env = dict(MASTER_ADDR=str(master_addr), MASTER_PORT=str(master_port), WORLD_SIZE=str(world_size),
                   CROSS_RANK=str(node_rank), CROSS_SIZE=str(len(node_list)), LOCAL_SIZE=str(len(local_gpu_ids)),
                   RANK=str(global_rank_mapping[local_node][local_rank]), LOCAL_RANK=str(local_rank),
                   CUDA_VISIBLE_DEVICES=",".join(map(str, local_gpu_ids)))
        envs.append(env)
    return envs, world_size


def main(args=None):
    args = parse_args(args)
    assert args.world_info != "None", "must provide world info dict"
    world_info = decode_world_info(args.world_info)
    envs, world_size = build_rank_env(world_info, args.node_rank, args.master_addr, args.master_port)
    logger.info(f"nnodes={len(world_info)}, node_rank={args.node_rank}, world_size={world_size}")
    processes = []
    log_dir = None if args.enable_each_rank_log == "None" else args.enable_each_rank_log
    if log_dir:
        os.makedirs(log_dir, exist_ok=True)
    for local_rank, rank_env in enumerate(envs):
        env = os.environ.copy()
        env.update(rank_env)
        cmd = []
        if not args.no_python:
            cmd = [sys.executable, "-u"]
            if args.module:
                cmd.append("-m")
        cmd.append(args.training_script)
        if not args.no_local_rank:
            cmd.append(f"--local_rank={local_rank}")
        cmd += args.training_script_args
        if args.bind_cores_to_rank:
            from ..utils.numa import get_numactl_cmd
            cores, numactl = get_numactl_cmd(args.bind_core_list, len(envs), local_rank)
            env["OMP_NUM_THREADS"] = str(cores)
            cmd = numactl + cmd
        if log_dir:
            f = open(os.path.join(log_dir, f"rank_{rank_env['RANK']}.log"), "w")
            processes.append(subprocess.Popen(cmd, env=env, stdout=f, stderr=f))
        else:
            processes.append(subprocess.Popen(cmd, env=env))
    if args.save_pid:
        with open(os.path.join(PID_FILE_BASEPATH, f"{os.getpid()}.deepspeed"), "w") as fd:
            fd.write(str(os.getpid()))

    def sigkill_handler(signum, frame):
        for p in processes:
            terminate_process_tree(p.pid)
        sys.exit(1)

    signal.signal(signal.SIGINT, sigkill_handler)
    signal.signal(signal.SIGTERM, sigkill_handler)
    alive = set(processes)
    while alive:
        finished = [p for p in alive if p.poll() is not None]
        for p in finished:
            if p.returncode != 0:
                logger.error(f"rank process {p.pid} exits with return code = {p.returncode}; killing the others")
                for q in processes:
                    if q is not p:
                        terminate_process_tree(q.pid)
                sys.exit(p.returncode)
            alive.remove(p)
        time.sleep(1)


if __name__ == "__main__":
    main()
