"""Ulysses SP: all-to-all layout switches invert each other, DistributedAttention == local full attention,
a sequence-parallel Llama matches the single-rank model (reference tests/unit/sequence_parallelism/test_ulysses.py)."""
import pytest
import torch

from common import run_distributed


def _layout_roundtrip():
    import torch.distributed as dist
    from deepspeed_b200.sequence import heads_to_seq, seq_to_heads
    torch.manual_seed(dist.get_rank())
    x = torch.randn(2, 6, 4, 8, requires_grad=True)
    h = seq_to_heads(x, dist.group.WORLD)
    assert h.shape == (2, 12, 2, 8)
    back = heads_to_seq(h, dist.group.WORLD)
    ok = torch.equal(back, x)
    back.sum().backward()
    return ok and torch.allclose(x.grad, torch.ones_like(x))


def test_layout_switch_roundtrip():
    assert all(run_distributed(_layout_roundtrip, 2))


def _dist_attention():
    import torch.distributed as dist
    import torch.nn.functional as F
    from deepspeed_b200.sequence import DistributedAttention
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    B, S, H, D = 2, 16, 4, 8
    q, k, v = torch.randn(B, S, H, D), torch.randn(B, S, 2, D), torch.randn(B, S, 2, D)

    def local_attn(q, k, v):  # [b, s, h, d]
        rep = q.shape[2] // k.shape[2]
        kk, vv = k.repeat_interleave(rep, 2), v.repeat_interleave(rep, 2)
        o = F.scaled_dot_product_attention(q.transpose(1, 2), kk.transpose(1, 2), vv.transpose(1, 2), is_causal=True)
        return o.transpose(1, 2)

    ref = local_attn(q, k, v)
    s = S // world
    sl = slice(rank * s, (rank + 1) * s)
    da = DistributedAttention(local_attn, dist.group.WORLD)
    out = da(q[:, sl].contiguous(), k[:, sl].contiguous(), v[:, sl].contiguous(), batch_dim_idx=0)
    return float((out - ref[:, sl]).abs().max())


def test_distributed_attention_matches_local():
    errs = run_distributed(_dist_attention, 2)
    assert max(errs) < 1e-5, errs


def _sp_llama(sp):
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    from deepspeed_b200.sequence import enable_ulysses
    torch.manual_seed(0)
    cfg = LlamaConfig.tiny(dtype=torch.float32, gradient_checkpointing=False)
    model = LlamaForCausalLM(cfg)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 1, "sequence_parallel_size": sp,
        "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "zero_optimization": {"stage": 1}})
    rank = ds.comm.get_rank()
    torch.manual_seed(5)
    ids = torch.randint(0, 512, (1, 32))
    tgt = torch.full_like(ids, -100)
    tgt[:, :-1] = ids[:, 1:]
    if sp > 1:
        enable_ulysses(model)
        s = 32 // sp
        sl = slice(rank * s, (rank + 1) * s)
        pos = torch.arange(32).unsqueeze(0)[:, sl]
        n_valid = int((tgt != -100).sum())
        # Ulysses convention: every rank normalises by the GLOBAL token count; ZeRO sums grads over the SP ranks
        loss = eng(ids[:, sl], labels=tgt[:, sl], position_ids=pos, shift_labels=False, loss_tokens=n_valid)
    else:
        loss = eng(ids, labels=tgt, shift_labels=False)
    eng.backward(loss)
    eng.step()
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param
    w = safe_get_full_fp32_param(model.model.layers[0].self_attn.qkv_proj.weight).clone()
    return float(loss.detach()), w


def test_sequence_parallel_llama_matches_single_rank():
    single = run_distributed(_sp_llama, 1, args=(1, ))[0]
    sp = run_distributed(_sp_llama, 2, args=(2, ))
    # the per-rank losses (each normalised by the global token count) add up to the global loss
    assert abs(sum(o[0] for o in sp) - single[0]) < 1e-4
    torch.testing.assert_close(sp[0][1], single[1], rtol=1e-4, atol=1e-5)


def test_tiled_mlp_and_tiled_loss_match_untiled():
    from deepspeed_b200.sequence.ulysses_sp import TiledFusedLogitsLoss, tiled_mlp_forward
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(16, 64), torch.nn.GELU(), torch.nn.Linear(64, 16))
    x = torch.randn(2, 40, 16, requires_grad=True)
    y = tiled_mlp_forward(mlp, x, shards=4)
    y.pow(2).sum().backward()
    gx, gw = x.grad.clone(), mlp[0].weight.grad.clone()
    x.grad = None
    mlp.zero_grad()
    mlp(x).pow(2).sum().backward()
    torch.testing.assert_close(gx, x.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(gw, mlp[0].weight.grad, rtol=1e-4, atol=1e-5)
    head = torch.nn.Linear(16, 50, bias=False)
    labels = torch.randint(0, 50, (2, 40))
    labels[0, :5] = -100

    def loss_fn(mod, xt, yt):
        return torch.nn.functional.cross_entropy(mod(xt).float(), yt, ignore_index=-100, reduction="sum")

    x2 = x.detach().clone().requires_grad_(True)
    tl = TiledFusedLogitsLoss.apply(loss_fn, head, x2, labels, None, 5, list(head.parameters()), "mean")
    tl.backward()
    x3 = x.detach().clone().requires_grad_(True)
    gw_t = head.weight.grad.clone()
    head.zero_grad()
    ref = torch.nn.functional.cross_entropy(head(x3).view(-1, 50).float(), labels.view(-1), ignore_index=-100)
    ref.backward()
    assert abs(float(tl) - float(ref)) < 1e-5
    torch.testing.assert_close(x2.grad, x3.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(gw_t, head.weight.grad, rtol=1e-4, atol=1e-6)


def test_fpdt_chunked_attention_matches_full():
    import torch.nn.functional as F
    from deepspeed_b200.sequence.fpdt_layer import chunked_causal_attention
    torch.manual_seed(0)
    q, k, v = torch.randn(1, 48, 4, 8), torch.randn(1, 48, 2, 8), torch.randn(1, 48, 2, 8)
    out = chunked_causal_attention(q, k, v, chunk_size=16)
    kk, vv = k.repeat_interleave(2, 2), v.repeat_interleave(2, 2)
    ref = F.scaled_dot_product_attention(q.transpose(1, 2), kk.transpose(1, 2), vv.transpose(1, 2),
                                         is_causal=True).transpose(1, 2)
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-5)


def _uneven_heads(world):
    """3 heads over 2 ranks: DistributedAttention must equal single-device attention, forward and backward."""
    import deepspeed_b200 as ds
    from deepspeed_b200.sequence.layer import (DistributedAttention, head_shard_sizes, heads_to_seq, seq_to_heads,
                                               uneven_heads_all2all)
    from deepspeed_b200.utils import groups
    ds.init_distributed()
    groups.initialize(sp_size=world, force=True)
    g = groups._get_sequence_parallel_group()
    rank = ds.comm.get_rank()
    torch.manual_seed(0)
    B, S, H, D = 2, 8, 3, 4
    q, k, v = (torch.randn(B, S, H, D) for _ in range(3))
    w = torch.randn(B, S, H, D)

    def attn(q, k, v):
        o = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                                                             is_causal=True)
        return o.transpose(1, 2)

    qr = q.clone().requires_grad_(True)
    ref = attn(qr, k, v)
    (ref * w).sum().backward()
    s = S // world
    sl = slice(rank * s, (rank + 1) * s)
    ql = q[:, sl].clone().requires_grad_(True)
    out = DistributedAttention(attn, g)(ql, k[:, sl].contiguous(), v[:, sl].contiguous(), batch_dim_idx=0)
    (out * w[:, sl]).sum().backward()
    sizes = head_shard_sizes(H, world)
    x = q[:, sl].contiguous()
    hx = uneven_heads_all2all(x, scatter_idx=2, gather_idx=1, batch_dim_idx=0, group=g)
    back = uneven_heads_all2all(hx, scatter_idx=1, gather_idx=2, batch_dim_idx=0, group=g)
    rt = heads_to_seq(seq_to_heads(x, g), g, num_heads=H)
    return (float((out - ref[:, sl]).abs().max()), float((ql.grad - qr.grad[:, sl]).abs().max()), sizes,
            tuple(hx.shape), bool(torch.equal(back, x)), bool(torch.equal(rt, x)))


def test_ulysses_uneven_heads():
    res = run_distributed(_uneven_heads, 2, args=(2, ))
    for r, (err, gerr, sizes, shape, ok1, ok2) in enumerate(res):
        assert err < 1e-5 and gerr < 1e-5 and sizes == [2, 1] and ok1 and ok2
        assert shape == (2, 8, sizes[r], 4)


def test_tiled_loss_and_autograd_compute_mlp_match_plain():
    """TiledLoss (uneven shards, uneven -100 distribution, an all-ignored shard) and AutogradComputeMLP equal the
    plain computation in value and gradient."""
    import torch
    import torch.nn.functional as F
    from deepspeed_b200.sequence.ulysses_sp import AutogradComputeMLP, TiledLoss

    def loss_fn(logits=None, labels=None, vocab_size=None, shift_labels=None):
        return F.cross_entropy(logits.float().reshape(-1, vocab_size), shift_labels.reshape(-1), ignore_index=-100)

    torch.manual_seed(0)
    logits = torch.randn(2, 23, 11, requires_grad=True)
    labels = torch.randint(0, 11, (2, 23))
    labels[:, :6] = -100  # the first shard of 4 (6 tokens each) is ignored entirely
    labels[0, 10:17] = -100
    want = loss_fn(logits=logits, vocab_size=11, shift_labels=labels)
    (gw, ) = torch.autograd.grad(want, logits)
    got = TiledLoss.apply(loss_fn, logits, 11, labels, 4)
    (gg, ) = torch.autograd.grad(got, logits)
    torch.testing.assert_close(got, want)
    torch.testing.assert_close(gg, gw, rtol=1e-5, atol=1e-7)
    none = TiledLoss.apply(loss_fn, logits, 11, torch.full((2, 23), -100), 3)
    assert float(none) == 0.0 and float(torch.autograd.grad(none, logits)[0].abs().sum()) == 0.0

    mlp = torch.nn.Sequential(torch.nn.Linear(8, 32), torch.nn.GELU(), torch.nn.Linear(32, 8))
    x = torch.randn(3, 5, 8, requires_grad=True)
    y0 = mlp(x)
    g0 = torch.autograd.grad(y0.square().sum(), [x] + list(mlp.parameters()))
    y1 = AutogradComputeMLP.apply(lambda m, t: m(t), mlp, x)
    y1.square().sum().backward()
    torch.testing.assert_close(y1, y0)
    for a, b in zip([x.grad] + [p.grad for p in mlp.parameters()], g0):
        torch.testing.assert_close(a, b)


def test_fpdt_helper_gradients():
    import torch
    from deepspeed_b200.sequence.fpdt_layer import apply_rotary_pos_emb_backward, bias_gelu, bias_gelu_back
    from deepspeed_b200.sequence.layer import apply_rotary_pos_emb
    torch.manual_seed(0)
    x = torch.randn(2, 5, 3, 12, dtype=torch.float64, requires_grad=True)
    ang = torch.randn(5, 1, 4, dtype=torch.float64)
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)  # rotate 8 of 12 channels
    g = torch.randn_like(x)
    (want, ) = torch.autograd.grad(apply_rotary_pos_emb(x, cos, sin), x, g)
    torch.testing.assert_close(apply_rotary_pos_emb_backward(g, cos, sin), want)
    y = torch.randn(7, dtype=torch.float64, requires_grad=True)
    gy = torch.randn(7, dtype=torch.float64)
    (w2, ) = torch.autograd.grad(bias_gelu(y), y, gy)
    torch.testing.assert_close(bias_gelu_back(gy, y.detach()), w2, rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(bias_gelu(y).detach(), torch.nn.functional.gelu(y.detach(), approximate="tanh"), rtol=1e-6,
                               atol=1e-7)
    from deepspeed_b200.monitor import check_comet_availability, check_tb_availability
    assert isinstance(check_tb_availability(), bool) and isinstance(check_comet_availability(), bool)
