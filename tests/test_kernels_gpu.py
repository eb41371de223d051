"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op
(reference strategy: tests/unit/ops/* compare kernel vs torch)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from deepspeed_b200.ops import kernels
    assert kernels.available()
    return kernels


def _dev():
    return torch.device("cuda", 0)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("n", [1, 1000, 65536 * 3 + 17, 10_000_001])
def test_sumsq(K, dtype, n):
    x = torch.randn(n, device=_dev(), dtype=dtype)
    got = K.sumsq(x)
    ref = x.double().pow(2).sum()
    assert abs(float(got) - float(ref)) <= 2e-4 * float(ref) + 1e-6
    got2 = K.sumsq(x)  # scratch ticket must have reset itself
    assert float(got2) == float(got)


def _adam_ref(p, g, m, v, lr, b1, b2, eps, wd, step, adamw):
    p, g, m, v = p.double(), g.double(), m.double(), v.double()
    if not adamw and wd:
        g = g + wd * p
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    upd = (m / (1 - b1**step)) / ((v / (1 - b2**step)).sqrt() + eps)
    if adamw and wd:
        upd = upd + wd * p
    return p - lr * upd, m, v


@pytest.mark.parametrize("adamw", [True, False])
@pytest.mark.parametrize("gdtype", [torch.float32, torch.bfloat16])
def test_fused_adam_matches_reference(K, adamw, gdtype):
    from deepspeed_b200.ops.optimizers import FusedAdam
    torch.manual_seed(0)
    shapes = [(1000, ), (333, 7), (70000, ), (8, )]
    params = [torch.randn(s, device=_dev()) for s in shapes]
    refs = [(p.clone(), torch.zeros_like(p), torch.zeros_like(p)) for p in params]
    opt = FusedAdam([{"params": params[:2], "weight_decay": 0.1}, {"params": params[2:], "weight_decay": 0.0}],
                    lr=1e-2, betas=(0.9, 0.95), eps=1e-8, adam_w_mode=adamw)
    lps = {p: torch.zeros_like(p, dtype=torch.bfloat16) for p in params}
    for step in range(1, 4):
        grads = {p: torch.randn_like(p).to(gdtype) for p in params}
        opt.step_flat(grads, lps, grad_scale=0.5)
        for i, p in enumerate(params):
            wd = 0.1 if i < 2 else 0.0
            rp, rm, rv = _adam_ref(refs[i][0], grads[p].float() * 0.5, refs[i][1], refs[i][2], 1e-2, 0.9, 0.95, 1e-8, wd,
                                   step, adamw)
            refs[i] = (rp.float(), rm.float(), rv.float())
            torch.testing.assert_close(p, refs[i][0], rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(opt.state[p]["exp_avg"], refs[i][1], rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(lps[p], p.to(torch.bfloat16), rtol=0, atol=0)


def test_fused_lion_sgd_adagrad_vs_cpu_math(K):
    from deepspeed_b200.ops.optimizers import FusedAdagrad, FusedLion, FusedSGD
    torch.manual_seed(1)
    for cls, kw in ((FusedLion, dict(lr=1e-3, betas=(0.9, 0.99), weight_decay=0.1)),
                    (FusedAdagrad, dict(lr=1e-2, eps=1e-10, weight_decay=0.01)),
                    (FusedSGD, dict(lr=1e-2, momentum=0.9, weight_decay=0.01))):
        p_gpu = torch.randn(5000, device=_dev())
        p_cpu = p_gpu.cpu().clone()
        o_gpu, o_cpu = cls([p_gpu], **kw), cls([p_cpu], **kw)
        for _ in range(3):
            g = torch.randn(5000)
            o_gpu.step_flat({p_gpu: g.to(_dev())}, None, 1.0)
            o_cpu.step_flat({p_cpu: g}, None, 1.0)
        torch.testing.assert_close(p_gpu.cpu(), p_cpu, rtol=1e-5, atol=1e-6)


def test_fused_lamb_vs_cpu_math(K):
    from deepspeed_b200.ops.optimizers import FusedLamb
    torch.manual_seed(2)
    p_gpu = torch.randn(40000, device=_dev())
    p_cpu = p_gpu.cpu().clone()
    o_gpu, o_cpu = FusedLamb([p_gpu], lr=1e-2, weight_decay=0.01), FusedLamb([p_cpu], lr=1e-2, weight_decay=0.01)
    for _ in range(3):
        g = torch.randn(40000)
        o_gpu.step_flat({p_gpu: g.to(_dev())}, None, 1.0)
        o_cpu.step_flat({p_cpu: g}, None, 1.0)
    torch.testing.assert_close(p_gpu.cpu(), p_cpu, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("cols", [64, 1024, 4096, 8192, 16384])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm_fwd_bwd(K, cols, with_res):
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    rows = 300
    x = torch.randn(rows, cols, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    r = torch.randn(rows, cols, device=_dev(), dtype=torch.bfloat16, requires_grad=True) if with_res else None
    w = (1 + 0.1 * torch.randn(cols, device=_dev())).to(torch.bfloat16).requires_grad_(True)
    out = fused.rms_norm(x, w, 1e-5, r)
    y, newres = out if with_res else (out, None)
    dy = torch.randn_like(y)
    dres = torch.randn_like(y) if with_res else None
    (y.float() * dy.float()).sum().backward() if not with_res else \
        ((y.float() * dy.float()).sum() + (newres.float() * dres.float()).sum()).backward()
    # fp32 reference
    xf = x.detach().float().requires_grad_(True)
    rf = r.detach().float().requires_grad_(True) if with_res else None
    wf = w.detach().float().requires_grad_(True)
    hin = xf + rf if with_res else xf
    if with_res:
        hin_q = hin.detach().to(torch.bfloat16).float() + (hin - hin.detach())  # bf16-rounded value, identity gradient
    else:
        hin_q = hin
    yr = hin_q * torch.rsqrt(hin_q.pow(2).mean(-1, keepdim=True) + 1e-5) * wf
    loss = (yr * dy.float()).sum() + ((hin * dres.float()).sum() if with_res else 0)
    loss.backward()
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(x.grad.float(), xf.grad, rtol=3e-2, atol=3e-2)
    torch.testing.assert_close(w.grad.float(), wf.grad, rtol=3e-2, atol=0.25)
    if with_res:
        torch.testing.assert_close(newres.float(), (xf + rf).detach(), rtol=1e-2, atol=1e-2)
        torch.testing.assert_close(r.grad.float(), rf.grad, rtol=3e-2, atol=3e-2)


@pytest.mark.parametrize("cols", [768, 4096])
def test_layernorm_fwd_bwd(K, cols):
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    rows = 257
    x = (torch.randn(rows, cols, device=_dev()) * 2 + 0.5).to(torch.bfloat16).requires_grad_(True)
    w = (1 + 0.1 * torch.randn(cols, device=_dev())).to(torch.bfloat16).requires_grad_(True)
    b = (0.1 * torch.randn(cols, device=_dev())).to(torch.bfloat16).requires_grad_(True)
    y = fused.layer_norm(x, w, b, 1e-5)
    dy = torch.randn_like(y)
    (y.float() * dy.float()).sum().backward()
    xf, wf, bf = (t.detach().float().requires_grad_(True) for t in (x, w, b))
    yr = torch.nn.functional.layer_norm(xf, (cols, ), wf, bf, 1e-5)
    (yr * dy.float()).sum().backward()
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(x.grad.float(), xf.grad, rtol=3e-2, atol=3e-2)
    torch.testing.assert_close(w.grad.float(), wf.grad, rtol=3e-2, atol=0.25)
    torch.testing.assert_close(b.grad.float(), bf.grad, rtol=3e-2, atol=0.25)


@pytest.mark.parametrize("with_pos", [False, True])
def test_rope_fwd_bwd(K, with_pos):
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    B, S, H, D = 2, 128, 4, 128
    x = torch.randn(B, S, H, D, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    cos, sin = fused.rope_tables(512, D, 10000.0, _dev())
    pos = torch.randint(0, 512, (B, S), device=_dev()) if with_pos else None
    y = fused.apply_rope(x, cos, sin, pos)
    dy = torch.randn_like(y)
    (y.float() * dy.float()).sum().backward()
    xf = x.detach().float().requires_grad_(True)
    yr = fused._rope_torch(xf, cos, sin, pos, S)
    (yr * dy.float()).sum().backward()
    torch.testing.assert_close(y.float(), yr, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(x.grad.float(), xf.grad, rtol=1e-2, atol=1e-2)


def test_swiglu_fwd_bwd(K):
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    gu = torch.randn(333, 2 * 1024, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    y = fused.swiglu(gu)
    dy = torch.randn_like(y)
    (y.float() * dy.float()).sum().backward()
    gf = gu.detach().float().requires_grad_(True)
    g, u = gf.chunk(2, -1)
    yr = torch.nn.functional.silu(g) * u
    (yr * dy.float()).sum().backward()
    torch.testing.assert_close(y.float(), yr, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(gu.grad.float(), gf.grad, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("vocab", [1000, 32000, 128256])
def test_cross_entropy_inplace_grad(K, vocab):
    torch.manual_seed(0)
    rows = 64
    logits = (torch.randn(rows, vocab, device=_dev()) * 3).to(torch.bfloat16)
    tgt = torch.randint(0, vocab, (rows, ), device=_dev())
    tgt[3] = -100
    lf = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lf, tgt, ignore_index=-100, reduction="none")
    (ref.sum() * 0.25).backward()
    work = logits.clone()
    loss = K.cross_entropy_inplace_grad(work, tgt, -100, 0.25)
    torch.testing.assert_close(loss, ref.detach(), rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(work.float(), lf.grad, rtol=2e-2, atol=2e-3)


def test_linear_cross_entropy_matches_torch(K):
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    N, H, V = 512, 256, 4096
    h = torch.randn(N, H, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    w = (torch.randn(V, H, device=_dev()) * 0.05).to(torch.bfloat16).requires_grad_(True)
    t = torch.randint(0, V, (N, ), device=_dev())
    t[:7] = -100
    loss = fused.linear_cross_entropy(h, w, t, chunk_tokens=200)
    loss.backward()
    hf, wf = h.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(hf @ wf.t(), t, ignore_index=-100)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 2e-2
    torch.testing.assert_close(h.grad.float(), hf.grad, rtol=5e-2, atol=2e-4)
    torch.testing.assert_close(w.grad.float(), wf.grad, rtol=5e-2, atol=2e-4)


def test_sumsq_not_disturbed_by_kernels_sharing_scratch(K):
    """Regression: norm backward writes partial sums into the shared scratch; sumsq's ticket word must be unaffected."""
    from deepspeed_b200.ops import fused
    x = torch.randn(600, 4096, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    w = torch.ones(4096, device=_dev(), dtype=torch.bfloat16, requires_grad=True)
    fused.rms_norm(x, w).float().sum().backward()
    t = torch.randn(1_000_003, device=_dev())
    for _ in range(2):
        got = float(K.sumsq(t))
        ref = float(t.double().pow(2).sum())
        assert abs(got - ref) <= 2e-4 * ref


@pytest.mark.parametrize("with_pos", [False, True])
def test_rope_qkv_inplace_fwd_bwd(K, with_pos):
    """Fused in-place RoPE on the q / k heads of a QKV projection vs the per-tensor fp32 reference."""
    from deepspeed_b200.ops import fused
    torch.manual_seed(0)
    B, S, Hq, Hkv, D = 2, 128, 4, 2, 128
    src = torch.randn(B, S, (Hq + 2 * Hkv) * D, device=_dev(), dtype=torch.bfloat16)
    cos, sin = fused.rope_tables(512, D, 10000.0, _dev())
    pos = torch.randint(0, 512, (B, S), device=_dev()) if with_pos else None
    leaf = src.clone().requires_grad_(True)
    qkv = leaf * 1.0  # a non-leaf the function may modify in place
    q, k, v = fused.apply_rope_qkv(qkv, cos, sin, pos, Hq, Hkv, D)
    dq, dk, dv = torch.randn_like(q), torch.randn_like(k), torch.randn_like(v)
    ((q.float() * dq.float()).sum() + (k.float() * dk.float()).sum() + (v.float() * dv.float()).sum()).backward()
    ref = src.float().clone().requires_grad_(True)
    rq, rk, rv = ref.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
    rq = fused._rope_torch(rq.reshape(B, S, Hq, D), cos, sin, pos, S)
    rk = fused._rope_torch(rk.reshape(B, S, Hkv, D), cos, sin, pos, S)
    rv = rv.reshape(B, S, Hkv, D)
    ((rq * dq.float()).sum() + (rk * dk.float()).sum() + (rv * dv.float()).sum()).backward()
    torch.testing.assert_close(q.float(), rq, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(k.float(), rk, rtol=1e-2, atol=1e-2)
    torch.testing.assert_close(v.float(), rv, rtol=0, atol=0)
    torch.testing.assert_close(leaf.grad.float(), ref.grad, rtol=1e-2, atol=2e-2)
