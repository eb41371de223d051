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
    torch.cuda.set_device(0)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", 0))
    step, _, _ = bench.build_ours(args, torch, 1)
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
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            agg[ev.name][0] += ev.device_time_total / 2e3
            agg[ev.name][1] += 1
    rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
    total = sum(v[0] for v in agg.values())
    out = {"total_kernel_ms_per_step": round(total, 2),
           "top": [{"kernel": k[:110], "ms_per_step": round(v[0], 3), "launches_per_step": v[1] // 2} for k, v in rows[:40]]}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
