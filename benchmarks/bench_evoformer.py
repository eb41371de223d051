#!/usr/bin/env python
"""DS4Sci_EvoformerAttention: native tcgen05 forward (+ GEMM backward from the saved LSE) vs the materialised-score
PyTorch baseline, on the shapes of the reference's own benchmark (tests/benchmarks/DS4Sci_EvoformerAttention_bench.py:
N = 256 MSA rows, 4 heads of 32, 256 residues, batch 1..4)."""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops.deepspeed4science.evoformer_attn import DS4Sci_EvoformerAttention  # noqa: E402


def baseline(q, k, v, biases, scale):
    a = torch.matmul(q.transpose(-2, -3), k.transpose(-2, -3).transpose(-1, -2)) * scale
    for b in biases:
        a = a + b
    return torch.matmul(torch.softmax(a, dim=-1), v.transpose(-2, -3)).transpose(-2, -3)


def timed(fn, iters=5):
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
    dev, dt = torch.device("cuda"), torch.bfloat16
    N, H, D, L = 256, 4, 32, 256
    rows = []
    for B in (1, 2, 4):
        mk = lambda *s, g=True: torch.randn(*s, device=dev, dtype=dt, requires_grad=g)  # noqa: E731
        q, k, v = mk(B, N, L, H, D), mk(B, N, L, H, D), mk(B, N, L, H, D)
        b1, b2 = mk(B, N, 1, 1, L, g=False), mk(B, 1, H, L, L)
        d_out = torch.rand(B, N, L, H, D, device=dev, dtype=dt)
        scale = 1 / math.sqrt(D)

        def fwd_ours():
            with torch.no_grad():
                return DS4Sci_EvoformerAttention(q, k, v, [b1, b2])

        def fwd_base():
            with torch.no_grad():
                return baseline(q, k, v, [b1, b2], scale)

        def fb(fn):
            def run():
                for t in (q, k, v, b2):
                    t.grad = None
                fn().backward(d_out)
            return run

        t_of, t_bf = timed(fwd_ours), timed(fwd_base)
        t_ofb = timed(fb(lambda: DS4Sci_EvoformerAttention(q, k, v, [b1, b2])))
        os.environ["DSB200_EVO_BWD"] = "sdpa"
        try:
            t_sdpa = timed(fb(lambda: DS4Sci_EvoformerAttention(q, k, v, [b1, b2])))
        except Exception as e:  # noqa: BLE001
            t_sdpa = float("nan")
            print("sdpa backward failed:", repr(e)[:200], flush=True)
        os.environ["DSB200_EVO_BWD"] = "gemm"
        t_bfb = timed(fb(lambda: baseline(q, k, v, [b1, b2], scale)))
        flops = 4.0 * B * N * H * L * L * D
        row = {"batch": B, "N": N, "L": L, "H": H, "D": D, "ours_fwd_ms": round(t_of, 3), "baseline_fwd_ms": round(t_bf, 3),
               "fwd_speedup": round(t_bf / t_of, 2), "ours_fwd_bwd_ms": round(t_ofb, 3),
               "baseline_fwd_bwd_ms": round(t_bfb, 3), "fwd_bwd_speedup": round(t_bfb / t_ofb, 2), "ours_fwd_bwd_sdpa_ms": round(t_sdpa, 3),
               "fwd_useful_tflops": round(flops / t_of / 1e9, 1)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    print(json.dumps({"results": rows}))


if __name__ == "__main__":
    main()
