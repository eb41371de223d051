#!/usr/bin/env python
"""Kernel-time breakdown of one training step of the headline config (torch.profiler, CUDA activities only).
Diagnostic only: numbers taken under the profiler are never reported as benchmark values."""
import argparse
import collections
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--act-ckpt", default="auto")
    a = ap.parse_args()
    sys.argv = ["bench.py", "--layers", str(a.layers), "--act-ckpt", a.act_ckpt]
    args = bench.parse()
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    for k, v in dict(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT="29519").items():
        os.environ.setdefault(k, v)
    world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", lr))
    step, _, _ = bench.build_ours(args, torch, world)
    ids = torch.randint(0, 128256, (1, args.seq), device="cuda")
    for _ in range(3):
        step(ids)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step(ids)
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0.0, 0])
    t_min, t_max = None, None
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            agg[ev.name][0] += ev.device_time_total / 2e3
            agg[ev.name][1] += 1
            tr = ev.time_range
            t_min = tr.start if t_min is None else min(t_min, tr.start)
            t_max = tr.end if t_max is None else max(t_max, tr.end)
    if rank != 0:
        torch.distributed.barrier()
        return
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    total = sum(v[0] for v in agg.values())
    out = {"world": world, "total_kernel_ms_per_step": round(total, 2),
           "span_ms_per_step": round((t_max - t_min) / 2e3, 2) if t_min is not None else None,
           "top": [{"kernel": k[:110], "ms_per_step": round(v[0], 3), "launches_per_step": v[1] // 2} for k, v in rows[:40]]}
    print(json.dumps(out, indent=1))
    if world > 1:
        torch.distributed.barrier()


if __name__ == "__main__":
    main()
