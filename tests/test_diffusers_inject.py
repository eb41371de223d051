"""diffusers attention / transformer-block replacement (reference ops/transformer/inference/diffusers_*.py and
module_inject generic_injection). diffusers is not installed here: the test carries the math of its
``Attention`` / ``BasicTransformerBlock`` (v0.2x: pre-LN, self attention, cross attention, GEGLU)."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from deepspeed_b200.ops.transformer.inference.diffusers_attention import DeepSpeedDiffusersAttention
from deepspeed_b200.ops.transformer.inference.diffusers_transformer_block import DeepSpeedDiffusersTransformerBlock
from deepspeed_b200.parallel.replace_module import generic_injection


class Attention(nn.Module):

    def __init__(self, dim, heads, ctx_dim=None):
        super().__init__()
        self.heads = heads
        self.to_q = nn.Linear(dim, dim, bias=False)
        self.to_k = nn.Linear(ctx_dim or dim, dim, bias=False)
        self.to_v = nn.Linear(ctx_dim or dim, dim, bias=False)
        self.to_out = nn.ModuleList([nn.Linear(dim, dim), nn.Dropout(0.0)])

    def forward(self, x, encoder_hidden_states=None, **_):
        c = x if encoder_hidden_states is None else encoder_hidden_states
        B, L, C = x.shape
        sp = lambda t: t.view(B, -1, self.heads, C // self.heads).transpose(1, 2)  # noqa: E731
        o = F.scaled_dot_product_attention(sp(self.to_q(x)), sp(self.to_k(c)), sp(self.to_v(c)))
        return self.to_out[1](self.to_out[0](o.transpose(1, 2).reshape(B, L, C)))


class GEGLU(nn.Module):

    def __init__(self, dim, inner):
        super().__init__()
        self.proj = nn.Linear(dim, inner * 2)

    def forward(self, x):
        h, gate = self.proj(x).chunk(2, dim=-1)
        return h * F.gelu(gate)


class FeedForward(nn.Module):

    def __init__(self, dim):
        super().__init__()
        self.net = nn.ModuleList([GEGLU(dim, dim * 4), nn.Dropout(0.0), nn.Linear(dim * 4, dim)])

    def forward(self, x):
        for m in self.net:
            x = m(x)
        return x


class BasicTransformerBlock(nn.Module):

    def __init__(self, dim=64, heads=4, ctx_dim=48):
        super().__init__()
        self.norm1, self.norm2, self.norm3 = nn.LayerNorm(dim), nn.LayerNorm(dim), nn.LayerNorm(dim)
        self.attn1 = Attention(dim, heads)
        self.attn2 = Attention(dim, heads, ctx_dim)
        self.ff = FeedForward(dim)

    def forward(self, x, encoder_hidden_states=None, **_):
        x = self.attn1(self.norm1(x)) + x
        x = self.attn2(self.norm2(x), encoder_hidden_states=encoder_hidden_states) + x
        return self.ff(self.norm3(x)) + x


class TinyUNet(nn.Module):

    def __init__(self):
        super().__init__()
        self.blocks = nn.ModuleList([BasicTransformerBlock(), BasicTransformerBlock()])

    def forward(self, x, ctx):
        for b in self.blocks:
            x = b(x, encoder_hidden_states=ctx)
        return x


class Pipe:

    def __init__(self):
        self.unet = TinyUNet()


def test_injected_pipeline_matches_original():
    torch.manual_seed(0)
    pipe = Pipe()
    x, ctx = torch.randn(2, 16, 64), torch.randn(2, 7, 48)
    with torch.no_grad():
        want = pipe.unet(x, ctx)
    generic_injection(pipe, dtype=torch.float32)
    assert pipe.ds_injected == {"attention": 4, "transformer_block": 2}
    blk = pipe.unet.blocks[0]
    assert isinstance(blk, DeepSpeedDiffusersTransformerBlock) and isinstance(blk.attn_1, DeepSpeedDiffusersAttention)
    assert blk.attn_1.do_out_bias is False and blk.attn_1.attn_qkvw.shape == (64, 192) and blk.attn_2.attn_qkvw is None
    assert blk.attn_2.attn_kw.shape == (48, 64)
    with torch.no_grad():
        got = pipe.unet(x, ctx)
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    generic_injection(pipe)  # idempotent
    assert pipe.ds_injected == {"attention": 0, "transformer_block": 0}


def test_attention_module_is_inference_only():
    torch.manual_seed(0)
    a = Attention(32, 4)
    from deepspeed_b200.ops.transformer.inference.config import DeepSpeedInferenceConfig
    ds = DeepSpeedDiffusersAttention.from_module(a, DeepSpeedInferenceConfig(hidden_size=32, heads=4,
                                                                             dtype=torch.float32))
    x = torch.randn(1, 5, 32)
    with torch.no_grad():
        torch.testing.assert_close(ds(x), a(x), rtol=1e-4, atol=1e-5)
    y = ds(x.clone().requires_grad_(True))
    with pytest.raises(RuntimeError, match="Inference mode"):
        y.sum().backward()


def test_int8_feed_forward_weights():
    torch.manual_seed(0)
    from deepspeed_b200.ops.transformer.inference.diffusers_2d_transformer import Diffusers2DTransformerConfig
    b = BasicTransformerBlock()
    x, ctx = torch.randn(1, 8, 64), torch.randn(1, 3, 48)
    with torch.no_grad():
        want = b(x, encoder_hidden_states=ctx)
        ds = DeepSpeedDiffusersTransformerBlock(b, Diffusers2DTransformerConfig(int8_quantization=True))
        assert ds.ff1_w.dtype == torch.int8
        got = ds(x, encoder_hidden_states=ctx)
    assert (got - want).abs().max() < 0.05 * want.abs().max()
