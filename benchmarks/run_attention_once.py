#!/usr/bin/env python
"""Launches the attention kernels a few times each (for ncu captures): tcgen05 flash-attention forward on the
Llama-3-8B shape and the paged decode kernel over a blocked KV cache."""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import attention as A  # noqa: E402
from deepspeed_b200.ops import kernels as K  # noqa: E402


def main():
    dev = torch.device("cuda")
    H, Hkv, D, S = 32, 8, 128, 4096
    qkv = torch.randn(1, S, (H + 2 * Hkv) * D, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv.split([H * D, Hkv * D, Hkv * D], dim=-1)
    q, k, v = q.view(1, S, H, D), k.view(1, S, Hkv, D), v.view(1, S, Hkv, D)
    for _ in range(3):
        A.flash_attention_forward(q, k, v, causal=True)
    n_seq, ctx, bs = 64, 4096, 128
    nb_seq = ctx // bs
    cache = torch.randn(n_seq * nb_seq, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(n_seq * nb_seq, device=dev).to(torch.int32).view(n_seq, nb_seq).contiguous()
    qd = torch.randn(n_seq, H, D, device=dev, dtype=torch.bfloat16)
    tok_seq = torch.arange(n_seq, dtype=torch.int32, device=dev)
    tok_pos = torch.full((n_seq, ), ctx - 1, dtype=torch.int32, device=dev)
    for _ in range(3):
        K.ragged_paged_attention(qd, cache, tok_seq, tok_pos, table, 1.0 / math.sqrt(D), ctx)
    torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
