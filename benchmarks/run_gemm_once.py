#!/usr/bin/env python
"""Tiny driver for ncu captures of the tcgen05 GEMM: a few launches of one Llama shape (gate_up by default)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import kernels as K  # noqa: E402

M, N, Kd = (int(x) for x in (sys.argv[1:4] if len(sys.argv) >= 4 else (4096, 28672, 4096)))
a = torch.randn(M, Kd, device="cuda", dtype=torch.bfloat16)
b = torch.randn(N, Kd, device="cuda", dtype=torch.bfloat16)
c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(6):
    K.gemm_bf16_nt(a, b, c)
torch.cuda.synchronize()
ref = (a[:256].float() @ b[:512].float().t())
assert (c[:256, :512].float() - ref).abs().max().item() < 0.02 * ref.abs().max().item() + 0.02
print("ok")
