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


@pytest.mark.parametrize("S,D,which", [(256, 32, "both"), (203, 32, "both"), (384, 64, "bias2"), (130, 128, "bias1"),
                                        (256, 16, "none")])
def test_bias_attention_matches_reference(S, D, which):
    """tcgen05 forward with additive biases (Evoformer form) vs an fp32 PyTorch restatement."""
    from deepspeed_b200.ops.attention import bias_attention_supported, flash_attention_bias_forward
    torch.manual_seed(0)
    dev = torch.device("cuda")
    B2, N, H = 2, 3, 4
    q, k, v = (torch.randn(B2 * N, S, H, D, device=dev, dtype=torch.bfloat16) for _ in range(3))
    assert bias_attention_supported(q)
    b1 = (torch.randn(B2 * N, S, device=dev) * 2).bfloat16() if which in ("both", "bias1") else None
    if b1 is not None:
        b1[:, -5:] = -1e4  # masked keys, as the MSA mask bias does
    b2 = torch.randn(B2, H, S, S, device=dev).bfloat16() if which in ("both", "bias2") else None
    o, lse = flash_attention_bias_forward(q, k, v, bias1=b1, bias2=b2, bias2_div=N)
    torch.cuda.synchronize()
    scores = torch.einsum("bihd,bjhd->bhij", q.float(), k.float()) / math.sqrt(D)
    if b1 is not None:
        scores = scores + b1.float()[:, None, None, :]
    if b2 is not None:
        scores = scores + b2.float().repeat_interleave(N, dim=0)
    want = torch.einsum("bhij,bjhd->bihd", scores.softmax(-1), v.float())
    assert o.shape == q.shape and o.dtype == q.dtype
    assert (o.float() - want).abs().max().item() < 3e-2
    assert (lse - torch.logsumexp(scores, -1)).abs().max().item() < 2e-2


def test_evoformer_attention_native_forward_and_grads():
    """DS4Sci_EvoformerAttention on the GPU (native forward, GEMM backward from the saved LSE) vs the fp32 math."""
    from deepspeed_b200.ops.deepspeed4science.evoformer_attn import DS4Sci_EvoformerAttention
    torch.manual_seed(1)
    dev = torch.device("cuda")
    B, N, L, H, D = 1, 4, 192, 4, 32
    mk = lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16).requires_grad_(True)  # noqa: E731
    q, k, v = mk(B, N, L, H, D), mk(B, N, L, H, D), mk(B, N, L, H, D)
    b1, b2 = mk(B, N, 1, 1, L), mk(B, 1, H, L, L)
    out = DS4Sci_EvoformerAttention(q, k, v, [b1, b2])
    w = torch.randn_like(out)
    (out.float() * w.float()).sum().backward()
    got = [t.grad.float().clone() for t in (q, k, v, b1, b2)]
    qf, kf, vf, b1f, b2f = (t.detach().float().requires_grad_(True) for t in (q, k, v, b1, b2))
    scores = torch.einsum("bnihd,bnjhd->bnhij", qf, kf) / math.sqrt(D) + b1f + b2f
    ref = torch.einsum("bnhij,bnjhd->bnihd", scores.softmax(-1), vf)
    (ref * w.float()).sum().backward()
    assert (out.float() - ref).abs().max().item() < 3e-2
    for g, t, name in zip(got, (qf, kf, vf, b1f, b2f), "q k v bias1 bias2".split()):
        denom = t.grad.abs().max().item() + 1e-6
        assert (g - t.grad).abs().max().item() / denom < 5e-2, name
