"""DeepNVMe perf tools (reference ``deepspeed/nvme``: ``ds_io`` single run, ``ds_nvme_tune`` sweep over block size /
queue depth / threads)."""
import argparse
import itertools
import json
import os
import time

import torch


def run_io(path, size_mb=256, block_size=1 << 20, queue_depth=8, threads=4, read=True, use_direct=True):
    from ..ops.aio import aio_handle
    h = aio_handle(block_size, queue_depth, False, True, threads, use_direct)
    n = size_mb * 2**20
    buf = torch.empty(n, dtype=torch.uint8, pin_memory=torch.cuda.is_available())
    if not os.path.exists(path) or os.path.getsize(path) < n:
        buf.random_(0, 255)
        h.sync_pwrite(buf, path)
    t0 = time.time()
    if read:
        h.async_pread(buf, path)
    else:
        h.async_pwrite(buf, path)
    h.wait()
    dt = time.time() - t0
    return dict(op="read" if read else "write", gbps=n / dt / 1e9, seconds=dt, block_size=block_size,
                queue_depth=queue_depth, threads=threads)


def ds_io_main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--path", required=True)
    p.add_argument("--size_mb", type=int, default=256)
    p.add_argument("--block_size", type=int, default=1 << 20)
    p.add_argument("--queue_depth", type=int, default=8)
    p.add_argument("--threads", type=int, default=4)
    p.add_argument("--write", action="store_true")
    a = p.parse_args(argv)
    print(json.dumps(run_io(a.path, a.size_mb, a.block_size, a.queue_depth, a.threads, read=not a.write)))


def sweep_main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--nvme_dir", required=True)
    p.add_argument("--size_mb", type=int, default=256)
    a = p.parse_args(argv)
    path = os.path.join(a.nvme_dir, "ds_nvme_tune.bin")
    best = {}
    for bs, qd, th in itertools.product((1 << 18, 1 << 20, 1 << 22), (4, 16), (1, 4, 8)):
        for read in (True, False):
            r = run_io(path, a.size_mb, bs, qd, th, read)
            k = r["op"]
            if k not in best or r["gbps"] > best[k]["gbps"]:
                best[k] = r
    print(json.dumps({"aio": {"block_size": best["read"]["block_size"], "queue_depth": best["read"]["queue_depth"],
                              "thread_count": best["read"]["threads"], "single_submit": False, "overlap_events": True},
                      "best": best}, indent=1))
