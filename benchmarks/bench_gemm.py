#!/usr/bin/env python
"""tcgen05 GEMM vs cuBLAS on the Llama-3-8B layer shapes. CUDA-event timing, L2 flushed between iterations."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import kernels as K  # noqa: E402


def timeit(fn, flush, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    assert K.available()
    peak = 1643.2
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            peak = json.load(f)["bf16_tflops"]
    except Exception:
        pass
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device="cuda")
    shapes = [("qkv", 4096, 6144, 4096), ("o_proj", 4096, 4096, 4096), ("gate_up", 4096, 28672, 4096),
              ("down", 4096, 4096, 14336), ("lm_head", 4096, 128256, 4096), ("square8k", 8192, 8192, 8192)]
    out = []
    for name, M, N, Kd in shapes:
        a = torch.randn(M, Kd, device="cuda", dtype=torch.bfloat16)
        b = torch.randn(N, Kd, device="cuda", dtype=torch.bfloat16)
        c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * Kd
        t_ours = timeit(lambda: K.gemm_bf16_nt(a, b, c), flush)
        t_cublas = timeit(lambda: torch.matmul(a, b.t(), out=c), flush)
        rec = dict(name=name, M=M, N=N, K=Kd, ours_ms=t_ours, cublas_ms=t_cublas, ours_tflops=fl / t_ours / 1e9,
                   cublas_tflops=fl / t_cublas / 1e9, ours_frac_of_measured_peak=fl / t_ours / 1e9 / peak,
                   ours_vs_cublas=t_cublas / t_ours)
        out.append(rec)
        print(json.dumps(rec))
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/gemm_bench.json", "w") as f:
        json.dump(dict(measured_peak_tflops=peak, results=out), f, indent=1)


if __name__ == "__main__":
    main()
