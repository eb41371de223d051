"""csrc/token_ops.cu against plain PyTorch: row gather/scatter, weighted scatter-add, row sort, random-LTD autograd."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from deepspeed_b200.ops import kernels
    assert kernels.available(), "libdsb200.so must load on a GPU box"
    return kernels


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
def test_gather_scatter_rows(K, dtype, idx_dtype):
    torch.manual_seed(0)
    src = torch.randn(4096, 1024, device="cuda").to(dtype)
    idx = torch.randperm(4096, device="cuda")[:1500].to(idx_dtype)
    out = K.gather_rows(src, idx)
    assert torch.equal(out, src[idx.long()])
    dst = torch.zeros_like(src)
    K.scatter_rows_(dst, idx, out)
    want = torch.zeros_like(src)
    want[idx.long()] = out
    assert torch.equal(dst, want)
    # out-of-range indices (capacity drops) are skipped
    idx2 = idx.clone()
    idx2[::7] = -1
    out2 = K.gather_rows(src, idx2, out=torch.zeros(1500, 1024, device="cuda", dtype=dtype))
    keep = idx2 >= 0
    assert torch.equal(out2[keep], src[idx2[keep].long()]) and out2[~keep].abs().sum() == 0


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_scatter_add_rows(K, dtype):
    torch.manual_seed(1)
    src = torch.randn(3000, 512, device="cuda").to(dtype)
    idx = torch.randint(0, 1000, (3000, ), device="cuda")
    w = torch.rand(3000, device="cuda")
    dst = torch.zeros(1000, 512, device="cuda", dtype=dtype)
    K.scatter_add_rows_(dst, idx, src, w)
    want = torch.zeros(1000, 512, device="cuda").index_add_(0, idx, src.float() * w[:, None])
    tol = 5e-2 if dtype == torch.bfloat16 else 1e-4
    torch.testing.assert_close(dst.float(), want, rtol=tol, atol=tol)
    perm = torch.randperm(1000, device="cuda")
    dst2 = torch.ones(1000, 512, device="cuda", dtype=dtype)
    K.scatter_add_rows_(dst2, perm, src[:1000], unique=True)
    want2 = torch.ones(1000, 512, device="cuda")
    want2[perm] += src[:1000].float()
    torch.testing.assert_close(dst2.float(), want2, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("n", [5, 64, 1000, 4096])
def test_sort_rows(K, n):
    x = torch.randint(0, 100000, (37, n), device="cuda", dtype=torch.int32)
    want = torch.sort(x, dim=1).values
    K.sort_rows_i32_(x)
    assert torch.equal(x, want)


def test_random_ltd_autograd_on_gpu(K):
    from deepspeed_b200.ops.random_ltd import GatherTokens, ScatterTokens, gpt_sample_tokens
    torch.manual_seed(0)
    B, S, H, R = 2, 64, 128, 16
    idx, _ = gpt_sample_tokens(R, S, B, layers=1, device="cuda")
    assert torch.all(idx[0, :, 1:] > idx[0, :, :-1])
    for batch_first in (True, False):
        x = torch.randn((B, S, H) if batch_first else (S, B, H), device="cuda", dtype=torch.bfloat16,
                        requires_grad=True)
        n0 = K.launch_count()
        full, part = GatherTokens.apply(x, idx[0], batch_first)
        out = ScatterTokens.apply(full, part * 2, idx[0], batch_first)
        out.float().sum().backward()
        assert K.launch_count() > n0
        li = idx[0].long()
        want = x.detach().clone()
        grad = torch.ones_like(want)
        for b in range(B):
            if batch_first:
                want[b, li[b]] = x.detach()[b, li[b]] * 2
                grad[b, li[b]] = 2
            else:
                want[li[b], b] = x.detach()[li[b], b] * 2
                grad[li[b], b] = 2
        assert torch.equal(out.detach(), want)
        assert torch.equal(x.grad, grad)
