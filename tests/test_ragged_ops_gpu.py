"""FastGen ragged kernels (csrc/ragged_ops.cu): RoPE + blocked-KV append and paged attention vs fp32 PyTorch."""
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_rope(x, cos, sin, pos, rd):
    rot, rest = x[..., :rd].float(), x[..., rd:].float()
    c, s = cos[pos.long()].unsqueeze(1), sin[pos.long()].unsqueeze(1)
    x1, x2 = rot[..., :rd // 2], rot[..., rd // 2:]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s, rest], dim=-1)


def _make_batch(seq_lens_seen_new, block_size, dev):
    """-> tok_seq, tok_pos, block_table (a shuffled block pool), per-sequence block lists"""
    n_seq = len(seq_lens_seen_new)
    need = [-(-(seen + new) // block_size) for seen, new in seq_lens_seen_new]
    perm = torch.randperm(sum(need) + 3).tolist()
    blocks, it = [], 0
    for n in need:
        blocks.append(perm[it:it + n])
        it += n
    table = torch.zeros(n_seq, max(need), dtype=torch.int32)
    for r, b in enumerate(blocks):
        table[r, :len(b)] = torch.tensor(b, dtype=torch.int32)
    tok_seq, tok_pos = [], []
    for r, (seen, new) in enumerate(seq_lens_seen_new):
        tok_seq += [r] * new
        tok_pos += list(range(seen, seen + new))
    return (torch.tensor(tok_seq, dtype=torch.int32, device=dev), torch.tensor(tok_pos, dtype=torch.int32, device=dev),
            table.to(dev), blocks, sum(need) + 3)


@pytest.mark.parametrize("D,Hq,Hkv,rd", [(128, 8, 2, 128), (64, 4, 4, 32), (128, 8, 1, 0), (64, 4, 2, 64)])
def test_kv_append_rope_and_paged_attention(D, Hq, Hkv, rd):
    from deepspeed_b200.ops import fused, kernels as K
    torch.manual_seed(0)
    dev = torch.device("cuda")
    bs = 16
    # (tokens already in the cache, new tokens): a long decode, a prefill chunk, a mid-sequence chunk, a fresh prompt
    spec = [(700, 1), (0, 37), (90, 5), (0, 1)]
    tok_seq, tok_pos, table, blocks, n_blocks = _make_batch(spec, bs, dev)
    T = tok_seq.numel()
    cache = torch.zeros(n_blocks, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    hist_k, hist_v = [], []
    cos, sin = fused.rope_tables(1024, rd if rd else D, 10000.0, dev)
    # history already in the cache (rotated keys)
    for r, (seen, new) in enumerate(spec):
        k = torch.randn(seen, Hkv, D, device=dev, dtype=torch.bfloat16)
        v = torch.randn(seen, Hkv, D, device=dev, dtype=torch.bfloat16)
        hist_k.append(k)
        hist_v.append(v)
        for p in range(seen):
            cache[blocks[r][p // bs], 0, p % bs] = k[p]
            cache[blocks[r][p // bs], 1, p % bs] = v[p]
    qkv = torch.randn(T, (Hq + 2 * Hkv) * D, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    q, k, v = q.view(T, Hq, D), k.view(T, Hkv, D), v.view(T, Hkv, D)
    q_ref = _ref_rope(q, cos, sin, tok_pos, rd) if rd else q.float()
    k_ref = _ref_rope(k, cos, sin, tok_pos, rd) if rd else k.float()
    K.ragged_kv_append_rope(q, k, v, cache, cos if rd else None, sin if rd else None, tok_seq, tok_pos, table, rd)
    torch.testing.assert_close(q.float(), q_ref, rtol=2e-2, atol=2e-2)
    # cache contents of the new tokens
    t = 0
    for r, (seen, new) in enumerate(spec):
        for i in range(new):
            p = seen + i
            got_k = cache[blocks[r][p // bs], 0, p % bs].float()
            got_v = cache[blocks[r][p // bs], 1, p % bs]
            torch.testing.assert_close(got_k, k_ref[t], rtol=2e-2, atol=2e-2)
            assert torch.equal(got_v, v[t])
            t += 1
    # ---- attention ----
    scale = 1.0 / math.sqrt(D)
    out = K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, scale, max(s + n for s, n in spec))
    torch.cuda.synchronize()
    t = 0
    for r, (seen, new) in enumerate(spec):
        keys = torch.cat([hist_k[r].float(), k_ref[t:t + new].to(torch.bfloat16).float()], 0)   # cache holds bf16
        vals = torch.cat([hist_v[r].float(), v[t:t + new].float()], 0)
        for i in range(new):
            n_keys = seen + i + 1
            qq = q[t + i].float()  # [Hq, D] (rotated, bf16-rounded)
            kk = keys[:n_keys].repeat_interleave(Hq // Hkv, dim=1)  # [n, Hq, D]
            vv = vals[:n_keys].repeat_interleave(Hq // Hkv, dim=1)
            s = torch.einsum("hd,nhd->hn", qq, kk) * scale
            o = torch.einsum("hn,nhd->hd", torch.softmax(s, -1), vv)
            err = (out[t + i].float() - o).abs().max().item()
            assert err < 2e-2, (r, i, err)
        t += new


def test_paged_attention_split_kv_long_context():
    """One decode token over 5000 cached keys: the key range is split over CTAs and merged."""
    from deepspeed_b200.ops import kernels as K
    torch.manual_seed(1)
    dev = torch.device("cuda")
    D, Hq, Hkv, bs, n = 128, 32, 8, 64, 5000
    nb = -(-n // bs)
    cache = torch.randn(nb, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    table = torch.randperm(nb, device=dev).to(torch.int32).view(1, nb)
    q = torch.randn(1, Hq, D, device=dev, dtype=torch.bfloat16)
    tok_seq = torch.zeros(1, dtype=torch.int32, device=dev)
    tok_pos = torch.full((1, ), n - 1, dtype=torch.int32, device=dev)
    out = K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, 1.0 / math.sqrt(D), n)
    kv = cache[table[0].long()]  # [nb, 2, bs, Hkv, D]
    keys = kv[:, 0].reshape(-1, Hkv, D)[:n].float().repeat_interleave(Hq // Hkv, 1)
    vals = kv[:, 1].reshape(-1, Hkv, D)[:n].float().repeat_interleave(Hq // Hkv, 1)
    s = torch.einsum("hd,nhd->hn", q[0].float(), keys) / math.sqrt(D)
    o = torch.einsum("hn,nhd->hd", torch.softmax(s, -1), vals)
    assert (out[0].float() - o).abs().max().item() < 1e-2


def test_ragged_engine_native_matches_fallback():
    """Greedy generation through the ragged engine: native kernels vs the per-sequence SDPA path."""
    from deepspeed_b200.inference.v2.engine_v2 import greedy_generate
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    from deepspeed_b200.ops import kernels as K
    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=512)
    model = LlamaForCausalLM(cfg, device=dev).eval()
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() >= 2:
                p.normal_(0.0, 0.05)
    ids = torch.randint(0, 512, (3, 70), device=dev)
    n0 = K.launch_count()
    os.environ["DSB200_RAGGED_NATIVE"] = "1"
    a = greedy_generate(model, ids, max_new_tokens=12)
    assert K.launch_count() > n0
    os.environ["DSB200_RAGGED_NATIVE"] = "0"
    try:
        b = greedy_generate(model, ids, max_new_tokens=12)
    finally:
        os.environ["DSB200_RAGGED_NATIVE"] = "1"
    # bf16 ties may flip a late token; the prompts and the first generated tokens must agree
    assert torch.equal(a[:, :74], b[:, :74]), (a[:, 70:], b[:, 70:])
    assert (a == b).float().mean().item() > 0.95


def test_paged_prefill_tcgen05_matches_simt_kernel():
    """Ragged prefill (two long chunks, one of them continuing a cached prefix) on the tcgen05 kernel with the block
    table vs the SIMT paged kernel; decode tokens of the same batch are served by the SIMT kernel through tok_row."""
    from deepspeed_b200.ops import attention as A, kernels as K
    torch.manual_seed(3)
    dev = torch.device("cuda")
    D, Hq, Hkv, bs = 128, 8, 2, 128
    spec = [(0, 300), (200, 1), (384, 257), (0, 130), (77, 3)]  # (cached, new)
    tok_seq, tok_pos, table, blocks, n_blocks = _make_batch(spec, bs, dev)
    T = tok_seq.numel()
    cache = torch.randn(n_blocks, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)  # history + this chunk's keys
    q = torch.randn(T, Hq, D, device=dev, dtype=torch.bfloat16)
    scale = 1.0 / math.sqrt(D)
    want = K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, scale, max(s + n for s, n in spec))
    assert A.paged_prefill_supported(q, cache)
    starts, it = [], 0
    for _, n in spec:
        starts.append(it)
        it += n
    long_rows = [r for r, (_, n) in enumerate(spec) if n >= 128]
    cu_q = torch.tensor([starts[r] for r in long_rows], dtype=torch.int32, device=dev)
    q_len = torch.tensor([spec[r][1] for r in long_rows], dtype=torch.int32, device=dev)
    kv_len = torch.tensor([sum(spec[r]) for r in long_rows], dtype=torch.int32, device=dev)
    out = torch.zeros(T, Hq, D, device=dev, dtype=torch.bfloat16)
    A.flash_attention_paged_prefill(q, cache, cu_q, q_len, kv_len, table[torch.tensor(long_rows, device=dev)].contiguous(),
                                    max(spec[r][1] for r in long_rows), scale, out=out)
    rest = [r for r in range(len(spec)) if r not in long_rows]
    rows = torch.tensor([starts[r] + i for r in rest for i in range(spec[r][1])], dtype=torch.int32, device=dev)
    rseq = torch.tensor([r for r in rest for _ in range(spec[r][1])], dtype=torch.int32, device=dev)
    rpos = torch.tensor([spec[r][0] + i for r in rest for i in range(spec[r][1])], dtype=torch.int32, device=dev)
    K.ragged_paged_attention(q, cache, rseq, rpos, table, scale, max(sum(spec[r]) for r in rest), out=out, tok_row=rows)
    torch.cuda.synchronize()
    assert (out.float() - want.float()).abs().max().item() < 2e-2


def test_ragged_engine_long_prompt_uses_tcgen05_prefill():
    from deepspeed_b200.inference.v2.engine_v2 import InferenceEngineV2, RaggedInferenceEngineConfig
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    dev = torch.device("cuda")
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=512, hidden_size=512, intermediate_size=1024, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=1024)
    model = LlamaForCausalLM(cfg, device=dev).eval()
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() >= 2:
                p.normal_(0.0, 0.05)
    # a long prompt (tcgen05 prefill), a short chunk (SIMT), another long one, and a single token (tcgen05 decode)
    prompts = [torch.randint(0, 512, (400, )), torch.randint(0, 512, (5, )), torch.randint(0, 512, (150, )),
               torch.randint(0, 512, (1, ))]
    logits = {}
    for bs in (128, 64):  # 128-token blocks: tcgen05 prefill for the long chunks; 64: SIMT kernel for everything
        eng = InferenceEngineV2(model, RaggedInferenceEngineConfig(kv_block_size=bs, num_kv_blocks=64,
                                                                   max_ragged_batch_size=1024))
        a = eng.put([0, 1, 2, 3], prompts)
        # all-decode batch on top of it
        b = eng.put([0, 1, 2, 3], [torch.tensor([7]), torch.tensor([8]), torch.tensor([9]), torch.tensor([3])])
        if bs == 128:
            assert eng._meta is not None
        logits[bs] = (a.float(), b.float())
    for x, y in zip(logits[128], logits[64]):
        assert (x - y).abs().max().item() < 0.15 * max(1.0, y.abs().max().item()), (x - y).abs().max()
        assert (x.argmax(-1) == y.argmax(-1)).all()


@pytest.mark.parametrize("Hq,Hkv,n_seq,ctx", [(32, 8, 5, 1000), (8, 8, 3, 260), (16, 2, 2, 4097), (32, 8, 70, 129)])
def test_flash_decode_paged_tcgen05_matches_simt(Hq, Hkv, n_seq, ctx):
    """Decode attention on the tcgen05 kernel (heads as tile rows, split-KV + merge) vs the SIMT paged kernel."""
    from deepspeed_b200.ops import attention as A, kernels as K
    torch.manual_seed(4)
    dev = torch.device("cuda")
    D, bs = 128, 128
    lens = [max(1, ctx - 37 * i) for i in range(n_seq)]
    spec = [(n - 1, 1) for n in lens]
    tok_seq, tok_pos, table, blocks, n_blocks = _make_batch(spec, bs, dev)
    cache = torch.randn(n_blocks, 2, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    q = torch.randn(n_seq, Hq, D, device=dev, dtype=torch.bfloat16)
    scale = 1.0 / math.sqrt(D)
    want = K.ragged_paged_attention(q, cache, tok_seq, tok_pos, table, scale, max(lens))
    kv_len = torch.tensor(lens, dtype=torch.int32, device=dev)
    for splits in (None, 1, 3):
        got = A.flash_decode_paged(q, cache, kv_len, table, max(lens), scale, splits=splits)
        torch.cuda.synchronize()
        err = (got.float() - want.float()).abs().max().item()
        assert err < 2e-2, (splits, err)
