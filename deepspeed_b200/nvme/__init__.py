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


BENCH_LOG_DIR = "_aio_bench_logs"
READ_LOG_DIR, WRITE_LOG_DIR = "read", "write"
DEFAULT_SWEEP_CONFIG = {"block_size": ["128K", "1M", "4M"], "queue_depth": [4, 16, 32], "io_parallel": [1, 4, 8],
                        "single_submit": [False], "overlap_events": [True]}


def _bytes(text):
    """"400M" / "128K" / "2G" / plain integers."""
    t = str(text).strip().upper()
    mult = {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30}.get(t[-1])
    return int(float(t[:-1]) * mult) if mult else int(t)


def parse_sweep_arguments(argv=None):
    """Command line of ``ds_nvme_tune`` (reference ``nvme/perf_run_sweep.py:66``)."""
    p = argparse.ArgumentParser()
    p.add_argument("--nvme_dir", nargs="+", required=True, help="writeable director(ies) on the NVMe device(s)")
    p.add_argument("--sweep_config", type=str, default=None, help="json file with the lists to sweep")
    p.add_argument("--no_read", action="store_true")
    p.add_argument("--no_write", action="store_true")
    p.add_argument("--io_size", type=str, default="400M")
    p.add_argument("--gpu", action="store_true", help="pinned buffers / device transfers")
    p.add_argument("--gds", action="store_true", help="sweep the GPUDirect-Storage handle (host bounce here)")
    p.add_argument("--flush_page_cache", action="store_true", help="needs sudo")
    p.add_argument("--log_dir", type=str, default=BENCH_LOG_DIR)
    p.add_argument("--loops", type=int, default=1)
    p.add_argument("--verbose", action="store_true")
    return p.parse_args(argv)


def sweep_main(args=None):
    """Run every combination of the sweep configuration for reads and writes, one JSON log line per run under
    ``<log_dir>/read`` and ``<log_dir>/write``; then print the best configuration (``generate_main``)."""
    if args is None or isinstance(args, (list, tuple)):
        args = parse_sweep_arguments(args)
    cfg = dict(DEFAULT_SWEEP_CONFIG)
    if args.sweep_config:
        cfg.update(json.load(open(args.sweep_config)))
    size_mb = max(1, _bytes(args.io_size) >> 20)
    ops = ([] if args.no_read else [True]) + ([] if args.no_write else [False])
    for sub in (READ_LOG_DIR, WRITE_LOG_DIR):
        os.makedirs(os.path.join(args.log_dir, sub), exist_ok=True)
    for read in ops:
        sub = READ_LOG_DIR if read else WRITE_LOG_DIR
        for bs, qd, th in itertools.product(cfg["block_size"], cfg["queue_depth"], cfg["io_parallel"]):
            best = None
            for _ in range(max(1, args.loops)):
                if args.flush_page_cache:
                    os.system("sync; echo 1 | sudo tee /proc/sys/vm/drop_caches > /dev/null")
                per_dir = [run_io(os.path.join(d, "ds_nvme_tune.bin"), size_mb, _bytes(bs), int(qd), int(th), read)
                           for d in args.nvme_dir]
                r = dict(per_dir[0], gbps=sum(x["gbps"] for x in per_dir))
                best = r if best is None or r["gbps"] > best["gbps"] else best
            name = f"{sub}_bs{bs}_qd{qd}_t{th}.txt"
            with open(os.path.join(args.log_dir, sub, name), "w") as f:
                f.write(json.dumps(best) + "\n")
            if args.verbose:
                print(name, f"{best['gbps']:.2f} GB/s")
    return generate_main(args.log_dir)


def generate_main(log_dir):
    """Read the sweep logs and print the ``aio`` ds_config section of the fastest read configuration (write speed
    breaks ties), as ``ds_nvme_tune`` does at the end of a sweep (reference ``nvme/perf_generate_param.py:86``)."""
    results = {}
    for sub in (READ_LOG_DIR, WRITE_LOG_DIR):
        d = os.path.join(log_dir, sub)
        for name in sorted(os.listdir(d)) if os.path.isdir(d) else []:
            with open(os.path.join(d, name)) as f:
                r = json.loads(f.readline())
            key = (r["block_size"], r["queue_depth"], r["threads"])
            results.setdefault(key, {})[sub] = r["gbps"]
    if not results:
        raise FileNotFoundError(f"no sweep logs under {log_dir}")
    key = max(results, key=lambda k: (results[k].get(READ_LOG_DIR, 0.0), results[k].get(WRITE_LOG_DIR, 0.0)))
    aio = {"block_size": key[0], "queue_depth": key[1], "intra_op_parallelism": key[2], "single_submit": False,
           "overlap_events": True, "use_gds": False}
    out = {"aio": aio, "read_GB/s": results[key].get(READ_LOG_DIR), "write_GB/s": results[key].get(WRITE_LOG_DIR)}
    print(json.dumps(out, indent=3))
    return out
