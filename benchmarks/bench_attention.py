#!/usr/bin/env python
"""tcgen05 flash-attention vs the framework SDPA (cuDNN) on the Llama-3-8B shape; CUDA-event timed."""
import argparse
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import attention as A  # noqa: E402


def timeit(fn, iters=20, warm=5):
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seq", type=int, nargs="+", default=[4096, 8192])
    ap.add_argument("--bwd", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda")
    H, Hkv, D = 32, 8, 128
    rows = []
    for S in a.seq:
        for causal in (True, False):
            qkv = torch.randn(1, S, (H + 2 * Hkv) * D, device=dev, dtype=torch.bfloat16)
            q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
            q, k, v = q.view(1, S, H, D), k.view(1, S, Hkv, D), v.view(1, S, Hkv, D)
            flops = 4.0 * S * S * D * H * (0.5 if causal else 1.0)
            ours = timeit(lambda: A.flash_attention_forward(q, k, v, causal=causal))
            qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
            lib = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal,
                                                                                  enable_gqa=True))
            o, _ = A.flash_attention_forward(q, k, v, causal=causal)
            o_lib = torch.nn.functional.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, enable_gqa=True)
            err = (o.float() - o_lib.transpose(1, 2).float()).abs().max().item()
            row = {"seq": S, "causal": causal, "ours_ms": round(ours, 4), "sdpa_ms": round(lib, 4),
                   "ours_tflops": round(flops / ours / 1e9, 1), "sdpa_tflops": round(flops / lib / 1e9, 1),
                   "max_abs_diff_vs_sdpa": err}
            if a.bwd:
                do = torch.randn(1, S, H, D, device=dev, dtype=torch.bfloat16)
                o, lse = A.flash_attention_forward(q, k, v, causal=causal)
                if hasattr(A, "_native_backward_available") and A._native_backward_available():
                    tb = timeit(lambda: A.flash_attention_backward(q, k, v, o, lse, do, causal=causal), iters=10)
                    row["ours_bwd_ms"] = round(tb, 4)
                    row["ours_bwd_tflops"] = round(2.5 * flops / tb / 1e9, 1)
                qg, kg, vg = (t.detach().clone().requires_grad_(True) for t in (qt, kt, vt))
                og = torch.nn.functional.scaled_dot_product_attention(qg, kg, vg, is_causal=causal, enable_gqa=True)
                dot = do.transpose(1, 2)
                tl = timeit(lambda: torch.autograd.grad(og, (qg, kg, vg), dot, retain_graph=True), iters=10)
                row["sdpa_bwd_ms"] = round(tl, 4)
                row["sdpa_bwd_tflops"] = round(2.5 * flops / tl / 1e9, 1)
            rows.append(row)
            print(json.dumps(row), flush=True)
    print(json.dumps({"results": rows}))


if __name__ == "__main__":
    main()
