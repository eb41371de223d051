"""Evoformer attention (two biases + bias gradients), DeepSpeedTransformerLayer vs a plain PyTorch BERT layer,
channels-last bias add. Mirrors reference tests/unit/ops/{deepspeed4science,transformer,spatial}."""
import pytest
import torch
import torch.nn.functional as F

from deepspeed_b200.ops.deepspeed4science import DS4Sci_EvoformerAttention
from deepspeed_b200.ops.spatial import nhwc_bias_add, nhwc_bias_add_add, nhwc_bias_add_bias_add
from deepspeed_b200.ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer


def _ref_evoformer(q, k, v, b1, b2):
    qh, kh, vh = (t.permute(0, 1, 3, 2, 4) for t in (q, k, v))
    s = qh @ kh.transpose(-1, -2) / q.shape[-1]**0.5
    if b1 is not None:
        s = s + b1
    if b2 is not None:
        s = s + b2
    return (torch.softmax(s, -1) @ vh).permute(0, 1, 3, 2, 4)


@pytest.mark.parametrize("use_b1,use_b2", [(True, True), (True, False), (False, False)])
def test_evoformer_attention_forward_backward(use_b1, use_b2):
    torch.manual_seed(0)
    B, N, L, H, D = 2, 3, 20, 4, 8
    q, k, v = (torch.randn(B, N, L, H, D, dtype=torch.float64, requires_grad=True) for _ in range(3))
    b1 = torch.randn(B, N, 1, 1, L, dtype=torch.float64, requires_grad=True) if use_b1 else None
    b2 = torch.randn(B, 1, H, L, L, dtype=torch.float64, requires_grad=True) if use_b2 else None
    biases = [b for b in (b1, b2) if b is not None] if not (b1 is None and b2 is not None) else [b1, b2]
    out = DS4Sci_EvoformerAttention(q, k, v, biases)
    ref = _ref_evoformer(q, k, v, b1, b2)
    torch.testing.assert_close(out, ref)
    g = torch.randn_like(out)
    inputs = [t for t in (q, k, v, b1, b2) if t is not None]
    got = torch.autograd.grad(out, inputs, g)
    want = torch.autograd.grad(ref, inputs, g)
    for a, b in zip(got, want):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-8)


def test_evoformer_chunking_is_transparent(monkeypatch):
    from deepspeed_b200.ops.deepspeed4science import evoformer_attn as E
    torch.manual_seed(1)
    q, k, v = (torch.randn(1, 6, 18, 2, 8) for _ in range(3))
    b1 = torch.randn(1, 6, 1, 1, 18)
    full = DS4Sci_EvoformerAttention(q, k, v, [b1])
    monkeypatch.setattr(E, "_WORKSPACE_ELEMS", 2 * 18 * 18 * 2)  # two MSA rows per chunk
    torch.testing.assert_close(DS4Sci_EvoformerAttention(q, k, v, [b1]), full)


@pytest.mark.parametrize("pre_ln", [True, False])
def test_transformer_layer_matches_plain_bert_layer(pre_ln):
    torch.manual_seed(0)
    cfg = DeepSpeedTransformerConfig(batch_size=2, hidden_size=64, intermediate_size=256, heads=4,
                                     attn_dropout_ratio=0.0, hidden_dropout_ratio=0.0, num_hidden_layers=2,
                                     initializer_range=0.02, pre_layer_norm=pre_ln, gelu_checkpoint=True,
                                     attn_dropout_checkpoint=True)
    layer = DeepSpeedTransformerLayer(cfg)
    for p in (layer.attn_qkvb, layer.attn_ob, layer.inter_b, layer.output_b, layer.attn_nb, layer.norm_b):
        p.data.normal_(0, 0.02)
    x = torch.randn(2, 10, 64, requires_grad=True)
    mask = torch.zeros(2, 1, 1, 10)
    mask[1, ..., 7:] = -10000.0

    def ref(x):
        def attn(h):
            qkv = F.linear(h, layer.attn_qkvw, layer.attn_qkvb).view(2, 10, 3, 4, 16).permute(2, 0, 3, 1, 4)
            s = qkv[0] @ qkv[1].transpose(-1, -2) / 4.0 + mask
            c = (torch.softmax(s, -1) @ qkv[2]).transpose(1, 2).reshape(2, 10, 64)
            return F.linear(c, layer.attn_ow, layer.attn_ob)

        def mlp(h):
            return F.linear(F.gelu(F.linear(h, layer.inter_w, layer.inter_b)), layer.output_w, layer.output_b)

        ln1 = lambda h: F.layer_norm(h, (64, ), layer.attn_nw, layer.attn_nb, 1e-12)  # noqa: E731
        ln2 = lambda h: F.layer_norm(h, (64, ), layer.norm_w, layer.norm_b, 1e-12)  # noqa: E731
        if pre_ln:
            h = x + attn(ln1(x))
            return h + mlp(ln2(h))
        h = ln1(x + attn(x))
        return ln2(h + mlp(h))

    layer.train()
    out = layer(x, attention_mask=mask)
    want = ref(x)
    torch.testing.assert_close(out, want, rtol=1e-4, atol=1e-5)
    g1, = torch.autograd.grad(out.sum(), x, retain_graph=True)
    g2, = torch.autograd.grad(want.sum(), x)
    torch.testing.assert_close(g1, g2, rtol=1e-4, atol=1e-5)
    sd = layer.state_dict()
    assert set(sd) == {"attn_qkvw", "attn_qkvb", "attn_ow", "attn_ob", "attn_nw", "attn_nb", "inter_w", "inter_b",
                       "output_w", "output_b", "norm_w", "norm_b"}


def test_nhwc_bias_add_cpu_semantics():
    x = torch.randn(2, 8, 4, 4).to(memory_format=torch.channels_last)
    o = torch.randn(2, 8, 4, 4).to(memory_format=torch.channels_last)
    b, ob = torch.randn(8), torch.randn(8)
    torch.testing.assert_close(nhwc_bias_add(x, b), x + b.view(1, 8, 1, 1))
    torch.testing.assert_close(nhwc_bias_add_add(x, b, o), x + b.view(1, 8, 1, 1) + o)
    torch.testing.assert_close(nhwc_bias_add_bias_add(x, b, o, ob), x + b.view(1, 8, 1, 1) + o + ob.view(1, 8, 1, 1))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_nhwc_bias_add_kernel(dtype):
    from deepspeed_b200.ops import kernels as K
    assert K.available()
    torch.manual_seed(0)
    x = torch.randn(4, 320, 32, 32, device="cuda", dtype=dtype).to(memory_format=torch.channels_last)
    o = torch.randn_like(x)
    b, ob = torch.randn(320, device="cuda", dtype=dtype), torch.randn(320, device="cuda", dtype=dtype)
    n0 = K.launch_count()
    got = nhwc_bias_add_bias_add(x, b, o, ob)
    assert K.launch_count() == n0 + 1
    want = (x.float() + b.float().view(1, -1, 1, 1) + o.float() + ob.float().view(1, -1, 1, 1)).to(dtype)
    torch.testing.assert_close(got, want, rtol=1e-2, atol=1e-2)
    x3 = torch.randn(2, 77, 768, device="cuda", dtype=dtype)
    b3 = torch.randn(768, device="cuda", dtype=dtype)
    torch.testing.assert_close(nhwc_bias_add(x3, b3), (x3.float() + b3.float()).to(dtype), rtol=1e-2, atol=1e-2)
