import os, sys, torch
sys.path.insert(0, os.getcwd())
from deepspeed_b200.ops import kernels as K
from deepspeed_b200.ops.fp_quantizer import FP_Quantize
dev = torch.device("cuda")
N, Kd = 28672, 4096
w = torch.randn(N, Kd, device=dev, dtype=torch.bfloat16) * 0.05
fq = FP_Quantize(group_size=512)
q = fq.quantize(w, q_bits=8)
payload = q.payload.view(torch.uint8)
x = torch.randn(1, Kd, device=dev, dtype=torch.bfloat16)
for _ in range(3):
    K.wo_gemm(x, payload, q.scale, N, Kd, 512, 8)
torch.cuda.synchronize()
print("done")
