"""tcgen05 flash-attention forward (csrc/attention_fwd.cu) against an fp32 PyTorch reference of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(q, k, v, causal, scale):
    B, Sq, H, D = q.shape
    Skv, Hkv = k.shape[1], k.shape[2]
    qf, kf, vf = q.float().transpose(1, 2), k.float().transpose(1, 2), v.float().transpose(1, 2)
    rep = H // Hkv
    kf, vf = kf.repeat_interleave(rep, dim=1), vf.repeat_interleave(rep, dim=1)
    s = torch.matmul(qf, kf.transpose(-1, -2)) * scale
    if causal:
        i = torch.arange(Sq, device=q.device)[:, None] + (Skv - Sq)
        j = torch.arange(Skv, device=q.device)[None, :]
        s = s.masked_fill(j > i, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    o = torch.matmul(torch.softmax(s, dim=-1), vf)
    return o.transpose(1, 2), lse


@pytest.mark.parametrize("B,S,H,Hkv,causal", [(1, 256, 4, 4, True), (2, 512, 8, 2, True), (1, 1024, 4, 1, False),
                                              (1, 384, 2, 2, True), (2, 200, 4, 2, False), (1, 2048, 8, 2, True)])
def test_flash_attn_fwd_matches_fp32_reference(B, S, H, Hkv, causal):
    from deepspeed_b200.ops import attention as A
    torch.manual_seed(0)
    dev = torch.device("cuda")
    # views of one fused QKV projection, exactly as the model produces them (strided rows)
    qkv = torch.randn(B, S, (H + 2 * Hkv) * 128, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv.split([H * 128, Hkv * 128, Hkv * 128], dim=-1)
    q, k, v = q.view(B, S, H, 128), k.view(B, S, Hkv, 128), v.view(B, S, Hkv, 128)
    assert A.supported(q, k, v)
    scale = 1.0 / math.sqrt(128)
    o, lse = A.flash_attention_forward(q, k, v, causal=causal, scale=scale)
    torch.cuda.synchronize()
    o_ref, lse_ref = _ref(q, k, v, causal, scale)
    assert torch.isfinite(o.float()).all()
    err = (o.float() - o_ref).abs().max().item()
    assert err < 2e-2, f"max abs err {err}"
    torch.testing.assert_close(lse, lse_ref, rtol=1e-3, atol=2e-3)


def test_flash_attn_fwd_large_scores_and_skv_ne_sq():
    """Row maxima that grow along the key axis (exercises the lazy rescale) and Skv != Sq (non-causal)."""
    from deepspeed_b200.ops import attention as A
    torch.manual_seed(1)
    dev = torch.device("cuda")
    B, Sq, Skv, H = 1, 256, 640, 2
    q = torch.randn(B, Sq, H, 128, device=dev, dtype=torch.bfloat16) * 3
    k = torch.randn(B, Skv, H, 128, device=dev, dtype=torch.bfloat16)
    k = k * torch.linspace(0.2, 4.0, Skv, device=dev).view(1, Skv, 1, 1).to(torch.bfloat16)
    v = torch.randn(B, Skv, H, 128, device=dev, dtype=torch.bfloat16)
    o, lse = A.flash_attention_forward(q, k, v, causal=False)
    o_ref, lse_ref = _ref(q, k, v, False, 1.0 / math.sqrt(128))
    assert (o.float() - o_ref).abs().max().item() < 3e-2
    torch.testing.assert_close(lse, lse_ref, rtol=1e-3, atol=5e-3)


def test_flash_attention_autograd_matches_sdpa():
    from deepspeed_b200.ops import attention as A
    torch.manual_seed(2)
    dev = torch.device("cuda")
    B, S, H, Hkv = 1, 512, 4, 2
    q = torch.randn(B, S, H, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(B, S, Hkv, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(B, S, Hkv, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    do = torch.randn(B, S, H, 128, device=dev, dtype=torch.bfloat16)
    o = A.flash_attention(q, k, v, causal=True)
    o.backward(do)
    got = [t.grad.float().clone() for t in (q, k, v)]
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    o_ref, _ = _ref(qf, kf, vf, True, 1.0 / math.sqrt(128))
    o_ref.backward(do.float())
    for g, w in zip(got, (qf.grad, kf.grad, vf.grad)):
        rel = (g - w).abs().max().item() / max(w.abs().max().item(), 1e-6)
        assert rel < 3e-2, rel


@pytest.mark.parametrize("B,S,H,Hkv,causal", [(1, 128, 2, 2, True), (1, 256, 4, 2, True), (2, 512, 8, 2, True),
                                              (1, 384, 4, 4, False), (1, 1024, 8, 1, True)])
def test_flash_attn_bwd_matches_fp32_reference(B, S, H, Hkv, causal):
    """tcgen05 backward kernels (dQ and dK/dV, csrc/attention_bwd.cu) against autograd through the fp32 reference."""
    from deepspeed_b200.ops import attention as A
    torch.manual_seed(0)
    dev = torch.device("cuda")
    qkv = torch.randn(B, S, (H + 2 * Hkv) * 128, device=dev, dtype=torch.bfloat16)
    q, k, v = qkv.split([H * 128, Hkv * 128, Hkv * 128], dim=-1)
    q, k, v = q.view(B, S, H, 128), k.view(B, S, Hkv, 128), v.view(B, S, Hkv, 128)
    do = torch.randn(B, S, H, 128, device=dev, dtype=torch.bfloat16)
    scale = 1.0 / math.sqrt(128)
    assert A._native_backward_available()
    o, lse = A.flash_attention_forward(q, k, v, causal=causal, scale=scale)
    dq, dk, dv = A.flash_attention_backward(q, k, v, o, lse, do, causal=causal, scale=scale)
    torch.cuda.synchronize()
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    o_ref, _ = _ref(qf, kf, vf, causal, scale)
    o_ref.backward(do.float())
    for name, g, w in (("dq", dq, qf.grad), ("dk", dk, kf.grad), ("dv", dv, vf.grad)):
        assert torch.isfinite(g.float()).all(), name
        err = (g.float() - w).abs().max().item()
        ref = w.abs().max().item()
        assert err <= 2e-2 * max(ref, 1.0), (name, err, ref)


def test_fpdt_chunked_attention_on_native_kernels_matches_full_attention():
    """FPDT blockwise causal attention (chunks merged through the differentiable LSE output of the native kernels):
    outputs and q/k/v gradients equal full causal attention."""
    from deepspeed_b200.ops import attention as A
    from deepspeed_b200.sequence.fpdt_layer import chunked_causal_attention
    torch.manual_seed(5)
    dev = torch.device("cuda")
    B, S, H, Hkv = 1, 1024, 4, 2
    q = torch.randn(B, S, H, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(B, S, Hkv, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(B, S, Hkv, 128, device=dev, dtype=torch.bfloat16, requires_grad=True)
    do = torch.randn(B, S, H, 128, device=dev, dtype=torch.bfloat16)
    assert A.lse_attention_supported(q[:, :256], k[:, :256], v[:, :256])
    out = chunked_causal_attention(q, k, v, chunk_size=256)
    (out.float() * do.float()).sum().backward()
    got = [t.grad.float().clone() for t in (q, k, v)]
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    o_ref, _ = _ref(qf, kf, vf, True, 1.0 / math.sqrt(128))
    o_ref.backward(do.float())
    assert (out.float() - o_ref).abs().max().item() < 3e-2
    for name, g, w in zip("qkv", got, (qf.grad, kf.grad, vf.grad)):
        err, ref = (g - w).abs().max().item(), w.abs().max().item()
        assert err <= 3e-2 * max(ref, 1.0), (name, err, ref)
