#!/usr/bin/env python
"""Kernel breakdown of one DS4Sci_EvoformerAttention forward + backward (torch profiler, top device ops)."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops.deepspeed4science.evoformer_attn import DS4Sci_EvoformerAttention  # noqa: E402

dev, dt = torch.device("cuda"), torch.bfloat16
B, N, H, D, L = 1, 256, 4, 32, 256
mk = lambda *s, g=True: torch.randn(*s, device=dev, dtype=dt, requires_grad=g)  # noqa: E731
q, k, v = mk(B, N, L, H, D), mk(B, N, L, H, D), mk(B, N, L, H, D)
b1, b2 = mk(B, N, 1, 1, L, g=False), mk(B, 1, H, L, L)
d_out = torch.rand(B, N, L, H, D, device=dev, dtype=dt)
for _ in range(2):
    DS4Sci_EvoformerAttention(q, k, v, [b1, b2]).backward(d_out)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        DS4Sci_EvoformerAttention(q, k, v, [b1, b2]).backward(d_out)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=60))
