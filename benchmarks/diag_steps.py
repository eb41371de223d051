#!/usr/bin/env python
"""Per-step diagnostics of the headline config: device time, host time, allocator activity (finds stalls)."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    sys.argv = ["bench.py"] + sys.argv[1:]
    args = bench.parse()
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    for k, v in dict(RANK="0", WORLD_SIZE="1", LOCAL_RANK="0", MASTER_ADDR="127.0.0.1", MASTER_PORT="29519").items():
        os.environ.setdefault(k, v)
    world, rank, lr = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", lr))
    step, _, _ = bench.build_ours(args, torch, world)
    n = args.steps
    ids = [torch.randint(0, 128256, (1, args.seq), device="cuda") for _ in range(n)]
    rows = []
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
    torch.cuda.synchronize()
    ev[0].record()
    keys = ["num_alloc_retries", "num_device_alloc", "num_device_free", "num_sync_all_streams"]
    for i in range(n):
        t0 = time.perf_counter()
        step(ids[i])
        t1 = time.perf_counter()
        ev[i + 1].record()
        st = torch.cuda.memory_stats()
        rows.append([round((t1 - t0) * 1e3, 1)] + [st.get(k, 0) for k in keys] +
                    [round(st["reserved_bytes.all.current"] / 2**30, 2)])
    torch.cuda.synchronize()
    for i in range(n):
        rows[i].insert(0, round(ev[i].elapsed_time(ev[i + 1]), 1))
    print(json.dumps({"rank": rank, "cols": ["dev_ms", "host_ms"] + keys + ["reserved_gib"], "rows": rows}), flush=True)
    torch.distributed.barrier()


if __name__ == "__main__":
    main()
