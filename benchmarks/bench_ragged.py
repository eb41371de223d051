#!/usr/bin/env python
"""Decode / prefill attention over the blocked KV cache (csrc/ragged_ops.cu, csrc/attention_fwd.cu paged mode):
achieved HBM bandwidth of the decode kernel vs the measured copy roofline, and the prefill kernel's throughput."""
import json
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepspeed_b200.ops import attention as A  # noqa: E402
from deepspeed_b200.ops import kernels as K  # noqa: E402


def timeit(fn, iters=30, warm=5):
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
    dev = torch.device("cuda")
    Hq, Hkv, D, bs = 32, 8, 128, 128
    rows = []
    peak = None
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    for n_seq, ctx in [(64, 4096), (256, 2048), (8, 16384), (1, 32768)]:
        nb_seq = -(-ctx // bs)
        nb = n_seq * nb_seq
        cache = torch.randn(nb, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
        table = torch.randperm(nb, device=dev).to(torch.int32).view(n_seq, nb_seq).contiguous()
        q = torch.randn(n_seq, Hq, D, device=dev, dtype=torch.bfloat16)
        tok_seq = torch.arange(n_seq, dtype=torch.int32, device=dev)
        tok_pos = torch.full((n_seq, ), ctx - 1, dtype=torch.int32, device=dev)
        out = torch.empty_like(q)
        scale = 1.0 / math.sqrt(D)
        ms = timeit(lambda: K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, scale, ctx, out=out))
        kv_len = torch.full((n_seq, ), ctx, dtype=torch.int32, device=dev)
        out2 = torch.empty_like(q)
        ms_tc = timeit(lambda: A.flash_decode_paged(q, cache, kv_len, table, ctx, scale, out=out2))
        diff = (out.float() - out2.float()).abs().max().item()
        kv_bytes = n_seq * ctx * Hkv * D * 2 * 2
        # library path: gather the blocks of every sequence, then SDPA
        def lib():
            kv = cache[table.long()]  # [n_seq, nb_seq, 2, bs, Hkv, D]
            kk = kv[:, :, 0].reshape(n_seq, -1, Hkv, D)[:, :ctx].transpose(1, 2)
            vv = kv[:, :, 1].reshape(n_seq, -1, Hkv, D)[:, :ctx].transpose(1, 2)
            return torch.nn.functional.scaled_dot_product_attention(q.unsqueeze(2), kk, vv, enable_gqa=True)
        ms_lib = timeit(lib, iters=10, warm=2)
        row = {"mode": "decode", "sequences": n_seq, "context": ctx, "simt_ms": round(ms, 4),
               "simt_kv_gb_per_s": round(kv_bytes / ms / 1e6, 1), "tcgen05_ms": round(ms_tc, 4),
               "tcgen05_kv_gb_per_s": round(kv_bytes / ms_tc / 1e6, 1),
               "tcgen05_of_copy_roofline": round(kv_bytes / ms_tc / 1e6 / peak, 3) if peak else None,
               "gather_plus_sdpa_ms": round(ms_lib, 4), "max_abs_diff": diff}
        rows.append(row)
        print(json.dumps(row), flush=True)
        del cache
    # ragged prefill on the tcgen05 kernel: 8 prompts of 2048 tokens, and 4 chunks of 512 tokens on a 3584-token prefix
    for n_seq, new, cached in [(8, 2048, 0), (4, 512, 3584)]:
        nb_seq = -(-(new + cached) // bs)
        nb = n_seq * nb_seq
        cache = torch.randn(nb, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
        table = torch.randperm(nb, device=dev).to(torch.int32).view(n_seq, nb_seq).contiguous()
        T = n_seq * new
        q = torch.randn(T, Hq, D, device=dev, dtype=torch.bfloat16)
        cu_q = torch.arange(0, T, new, dtype=torch.int32, device=dev)
        q_len = torch.full((n_seq, ), new, dtype=torch.int32, device=dev)
        kv_len = torch.full((n_seq, ), new + cached, dtype=torch.int32, device=dev)
        out = torch.empty_like(q)
        ms = timeit(lambda: A.flash_attention_paged_prefill(q, cache, cu_q, q_len, kv_len, table, new, out=out))
        flops = 4.0 * D * Hq * n_seq * (new * cached + new * new / 2)
        tok_seq = torch.arange(n_seq, dtype=torch.int32, device=dev).repeat_interleave(new)
        tok_pos = (torch.arange(new, dtype=torch.int32, device=dev) + cached).repeat(n_seq)
        ms_simt = timeit(lambda: K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, 1.0 / math.sqrt(D),
                                                          new + cached), iters=3, warm=1)
        row = {"mode": "prefill", "sequences": n_seq, "new_tokens": new, "cached_tokens": cached, "tcgen05_ms": round(ms, 4),
               "tcgen05_tflops": round(flops / ms / 1e9, 1), "simt_paged_ms": round(ms_simt, 3)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    print(json.dumps({"results": rows}))


if __name__ == "__main__":
    main()
