"""v1 inference transformer layer (ops/transformer/inference): equivalence with a HuggingFace GPT-2 block, cached
token-by-token decoding, rotary / GQA / gated-MLP variants against a plain PyTorch restatement, MoE inference layer.
Reference: model_implementations/transformers/ds_transformer.py, ops/transformer/inference/*."""
import math

import pytest
import torch
import torch.nn.functional as F

from deepspeed_b200.ops.transformer import (DeepSpeedInferenceConfig, DeepSpeedMoEInference,
                                            DeepSpeedMoEInferenceConfig, DeepSpeedTransformerInference)
from deepspeed_b200.utils.types import ActivationFuncType, NormType


def _fill(layer, seed=0):
    g = torch.Generator().manual_seed(seed)
    for name, p in layer.named_parameters():
        if p.dim() == 1 and ("nw" in name or "norm_w" in name):
            p.data.copy_(1.0 + 0.1 * torch.randn(p.shape, generator=g))
        else:
            p.data.copy_(0.05 * torch.randn(p.shape, generator=g))


def test_matches_hf_gpt2_block_and_decodes_incrementally():
    from transformers import GPT2Config
    from transformers.models.gpt2.modeling_gpt2 import GPT2Block
    torch.manual_seed(0)
    hc = GPT2Config(n_embd=64, n_head=4, n_layer=1, resid_pdrop=0.0, attn_pdrop=0.0, embd_pdrop=0.0)
    hc._attn_implementation = "eager"
    blk = GPT2Block(hc, layer_idx=0).eval()
    cfg = DeepSpeedInferenceConfig(hidden_size=64, heads=4, num_hidden_layers=1, layer_norm_eps=hc.layer_norm_epsilon,
                                   dtype=torch.float32, pre_layer_norm=True, triangular_masking=True,
                                   return_tuple=False)
    layer = DeepSpeedTransformerInference(cfg).eval()
    with torch.no_grad():
        layer.norm_w.copy_(blk.ln_1.weight)
        layer.norm_b.copy_(blk.ln_1.bias)
        layer.attention.attn_qkvw.copy_(blk.attn.c_attn.weight.t())  # Conv1D stores [in, out]
        layer.attention.attn_qkvb.copy_(blk.attn.c_attn.bias)
        layer.attention.attn_ow.copy_(blk.attn.c_proj.weight.t())
        layer.attention.attn_ob.copy_(blk.attn.c_proj.bias)
        layer.mlp.attn_nw.copy_(blk.ln_2.weight)
        layer.mlp.attn_nb.copy_(blk.ln_2.bias)
        layer.mlp.inter_w.copy_(blk.mlp.c_fc.weight.t())
        layer.mlp.inter_b.copy_(blk.mlp.c_fc.bias)
        layer.mlp.output_w.copy_(blk.mlp.c_proj.weight.t())
        layer.mlp.output_b.copy_(blk.mlp.c_proj.bias)
    x = torch.randn(2, 9, 64)
    with torch.no_grad():
        # the HF block applies no mask of its own (the model passes it): hand it the causal one
        causal = torch.zeros(9, 9).masked_fill(torch.triu(torch.ones(9, 9, dtype=torch.bool), 1), float("-inf"))
        want = blk(x, attention_mask=causal[None, None])
        want = want[0] if isinstance(want, tuple) else want
        got = layer(x)
        assert torch.allclose(got, want, atol=1e-5), (got - want).abs().max()
        # token by token through the internal cache
        DeepSpeedTransformerInference.reset_cache(layer)
        outs = [layer(x[:, :4], use_cache=True)[0]]
        for t in range(4, 9):
            outs.append(layer(x[:, t:t + 1], use_cache=True)[0])
        inc = torch.cat(outs, dim=1)
        assert torch.allclose(inc, want, atol=1e-5), (inc - want).abs().max()
        # explicit layer_past instead of the internal cache
        DeepSpeedTransformerInference.reset_cache(layer)
        o1, past = layer(x[:, :5], get_present=True)
        o2, past2 = layer(x[:, 5:], layer_past=past, get_present=True)
        assert past2[0].shape[2] == 9
        assert torch.allclose(torch.cat([o1, o2], dim=1), want, atol=1e-5)


def _plain_llama_like(layer, x, cfg):
    """Restatement in plain PyTorch: RMSNorm, half-rotation rotary, GQA, gated SiLU MLP, pre-norm residual stream."""
    a, m = layer.attention, layer.mlp
    b, s, h = x.shape
    hq, hk, d = cfg.heads, cfg.num_kv, h // cfg.heads

    def rms(t, w):
        return t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + cfg.epsilon) * w

    y = rms(x, layer.norm_w)
    qkv = F.linear(y, a.attn_qkvw, a.attn_qkvb)
    q, k, v = qkv.split([hq * d, hk * d, hk * d], dim=-1)
    q, k, v = (t.view(b, s, -1, d).transpose(1, 2) for t in (q, k, v))
    inv = 1.0 / (cfg.rope_theta**(torch.arange(0, d, 2).float() / d))
    ang = torch.arange(s).float()[:, None] * inv[None]
    cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)

    def rot(t):
        t1, t2 = t[..., :d // 2], t[..., d // 2:]
        return t * cos + torch.cat([-t2, t1], -1) * sin

    q, k = rot(q), rot(k)
    k, v = k.repeat_interleave(hq // hk, 1), v.repeat_interleave(hq // hk, 1)
    sc = (q @ k.transpose(-1, -2)) / math.sqrt(d)
    sc = sc.masked_fill(torch.triu(torch.ones(s, s, dtype=torch.bool), 1), float("-inf"))
    ctx = (sc.softmax(-1) @ v).transpose(1, 2).reshape(b, s, h)
    x = x + F.linear(ctx, a.attn_ow) + a.attn_ob
    y = rms(x, m.attn_nw)
    up, gate = F.linear(y, m.inter_w, m.inter_b).chunk(2, -1)
    return x + F.linear(up * F.silu(gate), m.output_w) + m.output_b


def test_rotary_gqa_rmsnorm_gated_variant():
    cfg = DeepSpeedInferenceConfig(hidden_size=64, intermediate_size=96, heads=8, num_kv=2, num_hidden_layers=1,
                                   layer_norm_eps=1e-6, dtype=torch.float32, norm_type=NormType.RMSNorm,
                                   rotary_dim=8, rotate_half=True, rotate_every_two=False,
                                   mlp_act_func_type=ActivationFuncType.GATED_SILU, return_tuple=False)
    layer = DeepSpeedTransformerInference(cfg).eval()
    _fill(layer)
    assert layer.attention.attn_qkvw.shape == ((8 + 4) * 8, 64) and layer.mlp.inter_w.shape == (192, 64)
    x = torch.randn(2, 7, 64)
    with torch.no_grad():
        got = layer(x)
        want = _plain_llama_like(layer, x, cfg)
    assert torch.allclose(got, want, atol=2e-5), (got - want).abs().max()


def test_padding_mask_forms_agree_and_return_conventions():
    cfg = DeepSpeedInferenceConfig(hidden_size=32, heads=4, num_hidden_layers=1, dtype=torch.float32,
                                   triangular_masking=False)
    layer = DeepSpeedTransformerInference(cfg).eval()
    _fill(layer, 1)
    x = torch.randn(1, 6, 32)
    keep = torch.tensor([[1, 1, 1, 1, 0, 0]])
    additive = torch.zeros(1, 1, 1, 6).masked_fill(keep[:, None, None, :] == 0, -1e4)
    with torch.no_grad():
        DeepSpeedTransformerInference.reset_cache(layer)
        a = layer(x, input_mask=keep)
        DeepSpeedTransformerInference.reset_cache(layer)
        b = layer(x, attention_mask=additive)
        DeepSpeedTransformerInference.reset_cache(layer)
        c = layer(x[:, :4])  # the padded keys must not matter for the real tokens
    assert isinstance(a, tuple) and len(a) == 2  # return_tuple: (output, mask)
    assert torch.allclose(a[0][:, :4], b[0][:, :4], atol=1e-5) and torch.allclose(a[0][:, :4], c[0], atol=1e-5)
    cfg.return_tuple, cfg.return_single_tuple = False, True
    DeepSpeedTransformerInference.reset_cache(layer)
    assert len(layer(x)) == 1
    with pytest.raises(ValueError):
        layer(x, layer_past=(x, x), past_key_value=(x, x))


def test_moe_inference_layer_top1_equals_routed_expert():
    cfg = DeepSpeedMoEInferenceConfig(hidden_size=32, heads=4, num_hidden_layers=1, moe_experts=4, global_experts=4,
                                      k=1, return_tuple=False)
    assert cfg.dtype == torch.float32
    layer = DeepSpeedMoEInference(cfg).eval()
    _fill(layer, 2)
    x = torch.randn(2, 5, 32)
    with torch.no_grad():
        out = layer(x)
        # recompute by hand: attention part, then each token through its arg-max expert scaled by the gate prob
        from deepspeed_b200.ops.transformer.inference.ds_transformer import _norm
        layer.attention.reset_cache()
        h = x + layer.attention(_norm(cfg, x, layer.norm_w, layer.norm_b), None, None, False)[0] + layer.attention.attn_ob
        m_in = _norm(cfg, h, layer.attn_nw, layer.attn_nb)
        probs = F.softmax(F.linear(m_in, layer.gate_w), -1)
        w, idx = probs.max(-1)
        want = h.clone()
        for e in range(4):
            sel = idx == e
            want[sel] += w[sel][:, None] * layer.experts[e](m_in[sel])
    assert torch.allclose(out, want, atol=1e-5), (out - want).abs().max()
    res = DeepSpeedMoEInference(DeepSpeedMoEInferenceConfig(hidden_size=32, heads=4, moe_experts=2, global_experts=2,
                                                            mlp_type="residual", return_tuple=True)).eval()
    _fill(res, 3)
    with torch.no_grad():
        assert res(x)[0].shape == x.shape
