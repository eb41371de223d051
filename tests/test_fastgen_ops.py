"""FastGen row kernels (csrc/fastgen_ops.cu) and the LoCo / fake-quant kernels (csrc/quant.cu): CPU semantics, and on a
GPU the native kernels against the same functions evaluated in fp32 PyTorch."""
import pytest
import torch
import torch.nn.functional as F

from deepspeed_b200.ops import fastgen as G
from deepspeed_b200.ops import quantizer as Q


def _reference_suite(dev, dtype):
    torch.manual_seed(0)
    out = {}
    table = torch.randn(97, 64, device=dev).to(dtype)
    ptab = torch.randn(40, 64, device=dev).to(dtype)
    ids = torch.randint(0, 97, (3, 11), device=dev)
    pos = torch.randint(0, 38, (3, 11), device=dev)
    out["embed"] = G.embed(ids, table, pos, ptab, pos_offset=2)
    out["embed_ref"] = (table.float()[ids] + ptab.float()[pos + 2])
    hidden = torch.randn(33, 64, device=dev).to(dtype)
    last = torch.tensor([4, 10, 32], device=dev)
    out["lg"], out["lg_ref"] = G.logits_gather(hidden, last), hidden.float()[last]
    x = torch.randn(7, 5, 96, device=dev).to(dtype)
    b = torch.randn(96, device=dev).to(dtype)
    for act in G.ACTS:
        out[f"ba_{act}"] = G.bias_act(x, b, act)
        out[f"ba_{act}_ref"] = G._torch_act(act)(x.float() + b.float())
        if act != "identity":
            out[f"ga_{act}"] = G.gated_act(x, b, act)
            g, u = (x.float() + b.float()).chunk(2, -1)
            out[f"ga_{act}_ref"] = G._torch_act(act)(g) * u
    out["ba_nobias"], out["ba_nobias_ref"] = G.bias_act(x, None, "relu"), F.relu(x.float())
    logits = torch.randn(50, 12, device=dev)
    s, e, c, slot = G.topk_gating(logits, 3, True, with_slots=True)
    p = torch.softmax(logits, -1)
    rs, re = p.topk(3, -1)
    out["tk_scores"], out["tk_scores_ref"] = s, rs / rs.sum(-1, keepdim=True)
    assert torch.equal(e.long().sort(-1).values, re.sort(-1).values)
    assert torch.equal(c.long(), torch.bincount(re.reshape(-1), minlength=12))
    # slots enumerate every expert's assignments exactly once
    for ex in range(12):
        got = slot[e == ex].long().sort().values
        assert torch.equal(got, torch.arange(int(c[ex]), device=dev))
    s2, _, _ = G.topk_gating(logits, 2, False)
    out["tk_raw"], out["tk_raw_ref"] = s2, p.topk(2, -1).values
    # scatter to expert-contiguous rows, identity "experts", weighted gather
    starts = torch.cumsum(c.long(), 0) - c.long()
    rows = (starts[e.long()] + slot.long()).to(torch.int32)
    tok = torch.randn(50, 64, device=dev).to(dtype)
    scattered = G.moe_scatter(tok, rows, int(c.sum()))
    assert torch.equal(scattered[rows[:, 1].long()], tok)
    rows_drop = rows.clone()
    rows_drop[::5, 0] = -1
    out["mg"] = G.moe_gather(scattered, rows_drop, s)
    w = s * (rows_drop >= 0)
    out["mg_ref"] = tok.float() * w.sum(-1, keepdim=True)
    return out


def _close(out, rtol, atol):
    for k, v in out.items():
        if k.endswith("_ref"):
            continue
        torch.testing.assert_close(v.float(), out[k + "_ref"].float(), rtol=rtol, atol=atol, msg=lambda m, k=k: f"{k}: {m}")


def test_fastgen_ops_cpu_semantics():
    _close(_reference_suite(torch.device("cpu"), torch.float32), 1e-5, 1e-6)


def test_loco_and_fake_quant_cpu_semantics():
    torch.manual_seed(0)
    x = torch.randn(4 * 256)
    err = torch.zeros(4 * 256)
    q, s = Q.loco_quantize(x, err, groups=4, bits=8, beta=0.8)
    deq = Q.dequantize(q, s, 4, 8, dtype=torch.float32)
    torch.testing.assert_close(deq + err, x, rtol=0, atol=1e-6)  # nothing is lost: value = dequant + error
    q2, s2 = Q.loco_quantize(x, err, groups=4, bits=8, beta=1.0)  # the second round carries the error forward
    two_round = (deq + Q.dequantize(q2, s2, 4, 8, dtype=torch.float32)) / 2
    assert (two_round - x).abs().mean() < (deq - x).abs().mean()
    y = Q.fake_quantize(x.clone(), groups=4, bits=4, symmetric=True)
    assert y.view(4, 256)[0].unique().numel() <= 16 and (y - x).abs().max() <= x.abs().max() / 8 + 1e-6
    ya = Q.fake_quantize(x.clone(), groups=4, bits=4, symmetric=False)
    assert ya.view(4, 256)[1].unique().numel() <= 16
    assert float(ya.max()) == pytest.approx(float(x.max()), abs=1e-5) and (ya - x).abs().mean() < (y - x).abs().mean() * 1.2
    sr = torch.stack([Q.fake_quantize(x.clone(), 4, 3, True, stochastic=True, seed=i) for i in range(200)]).mean(0)
    assert (sr - x).abs().mean() < (Q.fake_quantize(x.clone(), 4, 3, True) - x).abs().mean() * 0.5  # unbiased


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_fastgen_ops_native_kernels(dtype):
    assert G._native(torch.zeros(8, 64, device="cuda", dtype=dtype)), "fastgen kernels not loaded"
    tol = dict(rtol=2e-2, atol=2e-2) if dtype == torch.bfloat16 else dict(rtol=4e-3, atol=4e-3)
    _close(_reference_suite(torch.device("cuda"), dtype), **tol)


@pytest.mark.gpu
def test_loco_and_fake_quant_native_kernels():
    dev = torch.device("cuda")
    torch.manual_seed(0)
    assert hasattr(Q._lib(), "dsb200_loco_group_quant"), "quant kernels not loaded"
    for bits in (8, 4):
        x = torch.randn(16 * 512, device=dev).bfloat16()
        err = torch.randn(16 * 512, device=dev) * 0.01
        err_old, err_cpu, x_cpu = err.clone(), err.cpu().clone(), x.cpu()
        q, s = Q.loco_quantize(x, err, 16, bits, beta=0.8)
        qc, sc = Q.loco_quantize(x_cpu.float(), err_cpu, 16, bits, beta=0.8)
        torch.testing.assert_close(s.cpu(), sc, rtol=1e-5, atol=1e-7)
        deq = Q.dequantize(q.cpu(), s.cpu(), 16, bits, dtype=torch.float32).to(dev)  # exact fp32 dequantisation
        torch.testing.assert_close(deq + err, x.float() + 0.8 * err_old, rtol=0, atol=1e-5)  # value = dequant + new error
        torch.testing.assert_close(err.cpu(), err_cpu, rtol=0, atol=float(sc.max()) * 1.01)
        assert (q.cpu().view(torch.uint8).long() - qc.view(torch.uint8).long()).abs().float().mean() < 0.02
        qs, ss = Q.loco_quantize(x, torch.zeros_like(err), 16, bits, beta=0.0, swizzle=(2, 2))
        qr, sr = Q.quantize(x, 16, bits, True, swizzle=(2, 2))
        assert torch.equal(qs, qr) and torch.equal(ss, sr)
    for sym in (True, False):
        x = torch.randn(8 * 256, device=dev).bfloat16()
        got = Q.fake_quantize(x.clone(), 8, 4, sym)
        want = Q.fake_quantize(x.cpu().float().clone(), 8, 4, sym)
        assert (got.cpu().float() - want).abs().max() <= float(x.abs().max()) / 7 * 0.51 + 0.02
        assert got.view(8, 256)[0].float().unique().numel() <= 16
    x = torch.randn(8 * 256, device=dev).half()
    mean = torch.stack([Q.fake_quantize(x.clone(), 8, 3, True, True, seed=i).float() for i in range(200)]).mean(0)
    assert (mean - x.float()).abs().mean() < (Q.fake_quantize(x.clone(), 8, 3, True).float() - x.float()).abs().mean() * 0.5


@pytest.mark.gpu
def test_fastgen_decoder_blocks_use_the_kernels_and_match_the_loop_path(monkeypatch):
    """FastGen MoE block: gating kernel + scatter + grouped GEMM + gather against the per-expert loop; MLP blocks with
    bias / GeGLU through the fused activation kernels against plain PyTorch."""
    from deepspeed_b200.inference.v2.model_implementations import arch as A
    dev = torch.device("cuda")
    torch.manual_seed(0)
    spec = A.ArchSpec(hidden_size=128, intermediate_size=256, num_experts=8, num_experts_per_tok=2,
                      moe_intermediate_size=192, shared_expert_intermediate_size=64, dtype=torch.bfloat16)
    moe = A._MoE(spec, device=dev)
    for p in moe.parameters():
        torch.nn.init.normal_(p, std=0.05)
    x = torch.randn(3, 37, 128, device=dev, dtype=torch.bfloat16)
    assert moe._grouped_ok(x.view(-1, 128))
    with torch.no_grad():
        fast = moe(x)
        monkeypatch.setenv("DSB200_FASTGEN_GROUPED_MOE", "0")
        slow = moe(x)  # the experts' weights are views of the stacks now: same numbers through the loop
    torch.testing.assert_close(fast.float(), slow.float(), rtol=3e-2, atol=3e-2)
    assert moe.experts[3].up.weight.data_ptr() == moe._w13[3].data_ptr()
    for act, gated in (("gelu", True), ("gelu_tanh", False), ("relu", True)):
        sp = A.ArchSpec(hidden_size=128, intermediate_size=256, act=act, gated_mlp=gated, mlp_bias=True,
                        dtype=torch.bfloat16)
        mlp = A._MLP(sp, 256, device=dev)
        for p in mlp.parameters():
            torch.nn.init.normal_(p, std=0.05)
        with torch.no_grad():
            got = mlp(x)
            u = mlp.up(x).float()
            f = A._act(act)
            ref = (f(u[..., :256]) * u[..., 256:]) if gated else f(u)
            want = torch.nn.functional.linear(ref.bfloat16(), mlp.down.weight, mlp.down.bias)
        torch.testing.assert_close(got.float(), want.float(), rtol=3e-2, atol=3e-2)
