"""Block-sparse attention: layouts against loop-built expectations; SparseSelfAttention against dense attention with
the equivalent element mask. Mirrors reference tests/unit/ops/sparse_attention/test_sparse_attention.py."""
import pytest
import torch

from deepspeed_b200.ops.sparse_attention import (BertSparseSelfAttention, BigBirdSparsityConfig,
                                                 BSLongformerSparsityConfig, DenseSparsityConfig, FixedSparsityConfig,
                                                 LocalSlidingWindowSparsityConfig, SparseSelfAttention,
                                                 VariableSparsityConfig)


def _fixed_loops(num_heads, nb, L, G, attention, horizontal, patterns, per_head):
    layout = torch.zeros(num_heads, nb, nb, dtype=torch.int64)
    for h in range(num_heads if per_head else 1):
        for i in range(0, nb, L):
            end = min(i + L, nb)
            for row in range(i, end):
                for col in range(i, (row + 1 if attention == "unidirectional" else end)):
                    layout[h, row, col] = 1
        first = L - (1 + h % patterns) * G
        end = nb - (nb % L)
        for i in range(first, end, L):
            layout[h, (0 if attention == "bidirectional" else i):, i:i + G] = 1
            if horizontal:
                layout[h, i:i + G, :] = 1
        if end < nb:
            s = min(end + first, nb - G)
            layout[h, (0 if attention == "bidirectional" else s):, s:s + G] = 1
            if horizontal:
                layout[h, s:s + G, :] = 1
    if not per_head:
        layout[1:] = layout[0]
    return layout


@pytest.mark.parametrize("attention,horizontal", [("bidirectional", False), ("bidirectional", True),
                                                  ("unidirectional", False)])
def test_fixed_layout_matches_reference_algorithm(attention, horizontal):
    cfg = FixedSparsityConfig(num_heads=4, block=16, different_layout_per_head=True, num_local_blocks=4,
                              num_global_blocks=1, attention=attention, horizontal_global_attention=horizontal,
                              num_different_global_patterns=4)
    got = cfg.make_layout(16 * 10)
    want = _fixed_loops(4, 10, 4, 1, attention, horizontal, 4, True)
    assert torch.equal(got, want)
    assert not torch.equal(got[0], got[1])


def test_other_layouts_basic_properties():
    nb = 12
    dense = DenseSparsityConfig(num_heads=2, block=16).make_layout(16 * nb)
    assert dense.sum() == 2 * nb * nb
    var = VariableSparsityConfig(num_heads=2, block=16, local_window_blocks=[2, 4], global_block_indices=[0, 5],
                                 attention="unidirectional").make_layout(16 * nb)
    assert torch.equal(var[0], torch.tril(var[0])) and var[0, :, 0].all() and var[0, 5:, 5].all()
    assert var[0, 1, 0] == 1 and var[0, 2, 1] == 0 and var[0, 5, 2] == 1  # windows [0,2) [2,6) [6,10) ...
    bb = BigBirdSparsityConfig(num_heads=2, block=16, num_random_blocks=2, num_sliding_window_blocks=3,
                               num_global_blocks=1).make_layout(16 * nb)
    assert bb[0, 0].all() and bb[0, :, 0].all()
    assert all(bb[0, r, max(0, r - 1):r + 2].all() for r in range(nb))
    lf = BSLongformerSparsityConfig(num_heads=2, block=16, num_sliding_window_blocks=3,
                                    global_block_indices=[0, 7]).make_layout(16 * nb)
    assert lf[0, 7].all() and lf[0, :, 7].all() and lf[0, 3, 5] == 0
    sw = LocalSlidingWindowSparsityConfig(num_heads=2, block=16, num_sliding_window_blocks=3).make_layout(16 * nb)
    assert torch.equal(sw[0], torch.tril(sw[0])) and sw[0].sum() == nb + (nb - 1)
    with pytest.raises(ValueError):
        dense_cfg = DenseSparsityConfig(num_heads=2, block=16)
        dense_cfg.make_layout(100)


@pytest.mark.parametrize("cfg", [
    FixedSparsityConfig(num_heads=2, block=16, num_local_blocks=2, attention="bidirectional"),
    BigBirdSparsityConfig(num_heads=2, block=16, num_random_blocks=1, num_sliding_window_blocks=3),
    LocalSlidingWindowSparsityConfig(num_heads=2, block=16, num_sliding_window_blocks=3),
])
def test_sparse_self_attention_matches_masked_dense(cfg):
    torch.manual_seed(0)
    B, H, L, D = 2, 2, 128, 32
    attn = SparseSelfAttention(cfg, max_seq_length=256)
    q, k, v = (torch.randn(B, H, L, D) for _ in range(3))
    kpm = torch.zeros(B, L)
    kpm[1, -20:] = -10000.0
    rpe = torch.randn(H, L, L) * 0.1
    out = attn(q, k, v, rpe=rpe, key_padding_mask=kpm)
    layout = attn.get_layout(L)
    elem = layout.repeat_interleave(16, 1).repeat_interleave(16, 2).bool()     # [H, L, L]
    scores = q @ k.transpose(-1, -2) * D**-0.5 + rpe + kpm.view(B, 1, 1, L)
    scores = scores.masked_fill(~elem.unsqueeze(0), float("-inf"))
    ref = torch.softmax(scores, -1) @ v
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=1e-4)
    q.requires_grad_(True)
    attn(q, k, v).sum().backward()
    assert q.grad is not None and torch.isfinite(q.grad).all()


def test_bert_sparse_self_attention_shapes():
    from types import SimpleNamespace
    cfg = SimpleNamespace(hidden_size=64, num_attention_heads=4)
    layer = BertSparseSelfAttention(cfg, FixedSparsityConfig(num_heads=4, block=16, num_local_blocks=2))
    x = torch.randn(2, 64, 64)
    mask = torch.zeros(2, 64)
    assert layer(x, mask).shape == (2, 64, 64)


def test_block_sparse_matmul_modes_and_softmax_match_dense():
    """The standalone block-sparse ops (reference ops/sparse_attention/matmul.py, softmax.py) against dense math with
    the layout applied as a mask; an attention built from them equals SparseSelfAttention."""
    import torch
    from deepspeed_b200.ops.sparse_attention.matmul import MatMul, to_dense, to_sparse
    from deepspeed_b200.ops.sparse_attention.softmax import Softmax
    torch.manual_seed(0)
    B, H, L, D, blk = 2, 3, 64, 16, 16
    layout = (torch.rand(H, L // blk, L // blk) > 0.4).long()
    layout[:, torch.arange(L // blk), torch.arange(L // blk)] = 1  # every row keeps its diagonal block
    q, k, v = (torch.randn(B, H, L, D, requires_grad=True) for _ in range(3))
    mask = layout.repeat_interleave(blk, 1).repeat_interleave(blk, 2).bool()  # [H, L, L]
    # sdd: Q K^T on the pattern
    sdd = MatMul(layout, blk, "sdd", trans_b=True)
    w = sdd(q, k)
    assert w.shape == (B, int(layout.sum()), blk, blk)
    dense_w = torch.matmul(q, k.transpose(-1, -2))
    torch.testing.assert_close(w, to_sparse(dense_w, layout, blk), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(to_dense(w, layout, blk), dense_w * mask, rtol=1e-4, atol=1e-5)
    # softmax with scale, rpe, masks
    rpe = torch.randn(H, L, L)
    kpm = torch.zeros(B, L)
    kpm[:, -5:] = -1e4
    am = torch.tril(torch.ones(L, L))
    p = Softmax(layout, blk)(w, scale=D**-0.5, rpe=rpe, key_padding_mask=kpm, attn_mask=am,
                             key_padding_mask_mode="add", attn_mask_mode="mul")
    ref = dense_w * D**-0.5 + rpe + kpm[:, None, None, :] + (1 - am) * torch.finfo(torch.float32).min
    ref = torch.softmax(ref.masked_fill(~mask, float("-inf")), -1)
    torch.testing.assert_close(to_dense(p, layout, blk), ref, rtol=1e-4, atol=1e-6)
    # dsd: P V ; gradients flow through all three ops
    out = MatMul(layout, blk, "dsd")(p, v)
    want = torch.matmul(ref, v)
    torch.testing.assert_close(out, want, rtol=1e-4, atol=1e-5)
    g = torch.randn_like(out)
    got = torch.autograd.grad(out, (q, k, v), g, retain_graph=True)
    exp = torch.autograd.grad(want, (q, k, v), g)
    for a, b in zip(got, exp):
        torch.testing.assert_close(a, b, rtol=1e-3, atol=1e-5)
    # dds: dense x sparse, and the transposed variants
    x = torch.randn(B, H, 32, L)
    sp = to_sparse(dense_w.detach() * mask, layout, blk)
    torch.testing.assert_close(MatMul(layout, blk, "dds")(x, sp), torch.matmul(x, dense_w.detach() * mask), rtol=1e-4,
                               atol=1e-4)
    torch.testing.assert_close(MatMul(layout, blk, "dds", trans_b=True)(x, sp),
                               torch.matmul(x, (dense_w.detach() * mask).transpose(-1, -2)), rtol=1e-4, atol=1e-4)
    y = torch.randn(B, H, L, 8)
    torch.testing.assert_close(MatMul(layout, blk, "dsd", trans_a=True)(sp, y),
                               torch.matmul((dense_w.detach() * mask).transpose(-1, -2), y), rtol=1e-4, atol=1e-4)
    import pytest
    with pytest.raises(NotImplementedError):
        MatMul(layout, blk, "ddd")
