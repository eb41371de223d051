"""Fused tcgen05 GEMM + NVLink collective (csrc/gemm_comm.cu) vs the baseline of one NCCL collective plus one library
GEMM, on Llama-3-8B shapes. Device-timed (CUDA events), max over ranks, JSON on rank 0.

    torchrun --nproc-per-node N benchmarks/bench_gemm_comm.py [--tokens 4096] [--iters 20]
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
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
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from deepspeed_b200.comm import symm
    ctx = symm.SymmContext.get(dist.group.WORLD)
    dev = torch.device("cuda", torch.cuda.current_device())
    M, H, FF, QKV = args.tokens, 4096, 14336, 6144
    bf = torch.bfloat16
    res = {"world": world, "tokens_per_rank": M, "results": []}

    def record(name, shape, fused_ms, base_ms, flops, nvlink_bytes):
        res["results"].append({"op": name, "shape": shape, "fused_ms": round(fused_ms, 4), "nccl_plus_cublas_ms":
                               round(base_ms, 4), "speedup": round(base_ms / fused_ms, 3),
                               "fused_tflops": round(flops / fused_ms / 1e9, 1),
                               "nvlink_GBps_in_per_rank": round(nvlink_bytes / fused_ms / 1e6, 1)})

    # ---- (2) ZeRO-3 weight all-gather -> GEMM: gate_up projection, W [2*FF, H] row-sharded ----
    N = 2 * FF
    N = N // (world * 256) * (world * 256)
    x = torch.randn(M, H, device=dev).to(bf)
    wbuf = ctx.alloc(N // world * H * 2, want_multicast=False)
    wshard = wbuf.view(N // world * H, bf).view(N // world, H)
    wshard.normal_(0, 0.02)
    out = torch.empty(M, N, dtype=bf, device=dev)
    wfull = torch.empty(N, H, dtype=bf, device=dev)

    def base_ag():
        dist.all_gather_into_tensor(wfull, wshard)
        torch.matmul(x, wfull.t(), out=out)

    t_f = timed(lambda: ctx.all_gather_gemm(x, wbuf, N, out=out, gathered="b"), args.iters)
    t_b = timed(base_ag, args.iters)
    record("all_gather(W)->GEMM", [M, N, H], t_f, t_b, 2.0 * M * N * H, (world - 1) / world * N * H * 2)

    # ---- (4) Ulysses: QKV projection -> all-to-all ----
    Nq = QKV // (world * 128) * (world * 128)
    w = (torch.randn(Nq, H, device=dev) * 0.02).to(bf)
    obuf = ctx.alloc(world * M * (Nq // world) * 2, want_multicast=False)
    y = torch.empty(M, Nq, dtype=bf, device=dev)
    ysend = torch.empty(world, M, Nq // world, dtype=bf, device=dev)
    yrecv = torch.empty(world, M, Nq // world, dtype=bf, device=dev)

    def fused_a2a():
        ctx.gemm_all_to_all(x, w, obuf)
        ctx.release_a2a_out(obuf)

    def base_a2a():
        torch.matmul(x, w.t(), out=y)
        ysend.copy_(y.view(M, world, Nq // world).transpose(0, 1))
        dist.all_to_all_single(yrecv, ysend)

    t_f = timed(fused_a2a, args.iters)
    t_b = timed(base_a2a, args.iters)
    record("GEMM->all_to_all (QKV)", [M, Nq, H], t_f, t_b, 2.0 * M * Nq * H, (world - 1) / world * M * Nq * 2)

    # ---- AutoTP row-parallel: down projection partial -> reduce-scatter ----
    Kp = FF // world // 64 * 64
    Mr = M // (world * 128) * (world * 128)
    xp = torch.randn(Mr, Kp, device=dev).to(bf)
    wp = (torch.randn(H, Kp, device=dev) * 0.02).to(bf)
    sbuf = ctx.alloc(Mr * H * 2, want_multicast=False)
    o = torch.empty(Mr // world, H, dtype=bf, device=dev)
    part = torch.empty(Mr, H, dtype=bf, device=dev)

    def base_rs():
        torch.matmul(xp, wp.t(), out=part)
        dist.reduce_scatter_tensor(o, part)

    t_f = timed(lambda: ctx.gemm_reduce_scatter(xp, wp, sbuf, out=o), args.iters)
    t_b = timed(base_rs, args.iters)
    record("GEMM->reduce_scatter (down_proj TP)", [Mr, H, Kp], t_f, t_b, 2.0 * Mr * H * Kp,
           (world - 1) / world * Mr * H * 2)

    # ---- all-gather(X) -> GEMM (sequence-parallel input of a column-parallel linear) ----
    Ms = M // world // 128 * 128
    abuf = ctx.alloc(Ms * H * 2, want_multicast=False)
    ash = abuf.view(Ms * H, bf).view(Ms, H)
    ash.normal_()
    wc = (torch.randn(FF // world // 256 * 256, H, device=dev) * 0.02).to(bf)
    oc = torch.empty(world * Ms, wc.shape[0], dtype=bf, device=dev)
    xg = torch.empty(world * Ms, H, dtype=bf, device=dev)

    def base_aga():
        dist.all_gather_into_tensor(xg, ash)
        torch.matmul(xg, wc.t(), out=oc)

    t_f = timed(lambda: ctx.all_gather_gemm(wc, abuf, world * Ms, out=oc, gathered="a"), args.iters)
    t_b = timed(base_aga, args.iters)
    record("all_gather(X)->GEMM (TP column)", [world * Ms, wc.shape[0], H], t_f, t_b,
           2.0 * world * Ms * wc.shape[0] * H, (world - 1) / world * world * Ms * H * 2)
    if rank == 0:
        print(json.dumps(res))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
