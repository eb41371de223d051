#!/usr/bin/env python
"""Micro-benchmark: fused NVLink collectives (csrc/symm.cu) vs NCCL. Device-timed (CUDA events), max over ranks.

Launch: ``python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/bench_symm.py``.
Roofline denominator: measured peer copy 770 GB/s per direction per GPU (B200_PROFILING.md).
Bytes on the wire per rank: reduce-scatter / all-gather / all-to-all move (P-1)/P * N bytes in (and out).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LINK_GBS = 770.0


def timeit(fn, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from deepspeed_b200.comm import symm
    rank, world = dist.get_rank(), dist.get_world_size()
    ctx = symm.SymmContext.get(dist.group.WORLD)
    dev = torch.device("cuda", local)
    results = []
    for mib in (16, 128, 512):
        n = mib * 2**20 // 2  # bf16 elements of the full bucket
        n = n // (world * 128) * world * 128
        shard = n // world
        wire = (world - 1) / world * n * 2
        buf = ctx.alloc(n * 2)
        buf.view(n, torch.bfloat16).normal_()
        out = torch.zeros(shard, dtype=torch.bfloat16, device=dev)
        x = buf.view(n, torch.bfloat16)
        nccl_out = torch.empty(shard, dtype=torch.bfloat16, device=dev)
        for ctas in (16, 32, 64):
            for mc in ((False, True) if buf.mc_ptr else (False, )):
                ms = timeit(lambda: ctx.reduce_scatter(buf, torch.bfloat16, shard, out, False, 1.0 / world,
                                                       use_multicast=mc, ctas=ctas))
                results.append(dict(op="reduce_scatter", impl="symm_mc" if mc else "symm_p2p", ctas=ctas, mib=mib, ms=ms,
                                    busbw_gbs=wire / ms / 1e6, frac_of_link=wire / ms / 1e6 / LINK_GBS))
        ms = timeit(lambda: dist.reduce_scatter_tensor(nccl_out, x, op=dist.ReduceOp.AVG))
        results.append(dict(op="reduce_scatter", impl="nccl", mib=mib, ms=ms, busbw_gbs=wire / ms / 1e6,
                            frac_of_link=wire / ms / 1e6 / LINK_GBS))
        sb = ctx.alloc(shard * 2, want_multicast=False)
        full = torch.empty(n, dtype=torch.bfloat16, device=dev)
        for ctas in (16, 32, 64):
            ms = timeit(lambda: ctx.all_gather(sb, 0, shard * 2, full, ctas=ctas))
            results.append(dict(op="all_gather", impl="symm_p2p", ctas=ctas, mib=mib, ms=ms, busbw_gbs=wire / ms / 1e6,
                                frac_of_link=wire / ms / 1e6 / LINK_GBS))
        mine = sb.view(shard, torch.bfloat16)
        ms = timeit(lambda: dist.all_gather_into_tensor(full, mine))
        results.append(dict(op="all_gather", impl="nccl", mib=mib, ms=ms, busbw_gbs=wire / ms / 1e6,
                            frac_of_link=wire / ms / 1e6 / LINK_GBS))
        rows = 1024
        rowb = (n * 2 // world // rows) // 16 * 16
        ib = ctx.alloc(world * rows * rowb, want_multicast=False)
        o = torch.empty(world * rows * rowb, dtype=torch.uint8, device=dev)
        wire_a2a = (world - 1) * rows * rowb
        for ctas in (16, 32, 64):
            ms = timeit(lambda: ctx.all_to_all(ib, rows, rowb, o, ctas=ctas))
            results.append(dict(op="all_to_all", impl="symm_p2p", ctas=ctas, mib=mib, ms=ms,
                                busbw_gbs=wire_a2a / ms / 1e6, frac_of_link=wire_a2a / ms / 1e6 / LINK_GBS))
        src = ib.tensor[:world * rows * rowb]
        ms = timeit(lambda: dist.all_to_all_single(o, src))
        results.append(dict(op="all_to_all", impl="nccl", mib=mib, ms=ms, busbw_gbs=wire_a2a / ms / 1e6,
                            frac_of_link=wire_a2a / ms / 1e6 / LINK_GBS))
        del buf, sb, ib
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/symm_bench_n{world}.json", "w") as f:
            json.dump(dict(world=world, link_gbs=LINK_GBS, results=results), f, indent=1)
        for r in results:
            print(json.dumps(r))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
