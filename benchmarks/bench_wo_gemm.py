#!/usr/bin/env python
"""Weight-only quantized decode GEMM (csrc/wo_gemm.cu) vs the bf16 library GEMM on Llama-3-8B layer shapes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import kernels as K  # noqa: E402
from deepspeed_b200.ops.fp_quantizer import FP_Quantize  # noqa: E402


def timeit(fn, iters=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


BITS = tuple(int(b) for b in os.environ.get("WO_BITS", "8,6,4").split(","))


def main():
    dev = torch.device("cuda")
    rows = []
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)  # > L2: evict the weights between iterations
    for (N, Kd) in [(6144, 4096), (28672, 4096), (4096, 14336)]:
        w = torch.randn(N, Kd, device=dev, dtype=torch.bfloat16) * 0.05
        for bits in BITS:
            fq = FP_Quantize(group_size=512)
            q = fq.quantize(w, q_bits=bits)
            payload = q.payload.view(torch.uint8)
            for M in (1, 8, 32, 64):
                x = torch.randn(M, Kd, device=dev, dtype=torch.bfloat16)

                def ours():
                    flush.zero_()
                    return K.wo_gemm(x, payload, q.scale, N, Kd, 512, bits)

                def lib():
                    flush.zero_()
                    return torch.nn.functional.linear(x, w)

                t_flush = timeit(lambda: flush.zero_())
                t_ours, t_lib = timeit(ours) - t_flush, timeit(lib) - t_flush
                wbytes = payload.numel() + q.scale.numel() * 4
                row = {"N": N, "K": Kd, "bits": bits, "M": M, "wo_gemm_tc_us": round(t_ours * 1e3, 1),
                       "bf16_linear_us": round(t_lib * 1e3, 1), "weight_gb_per_s": round(wbytes / t_ours / 1e6, 1),
                       "speedup_vs_bf16": round(t_lib / t_ours, 2)}
                rows.append(row)
                print(json.dumps(row), flush=True)
    print(json.dumps({"results": rows}))


if __name__ == "__main__":
    main()
