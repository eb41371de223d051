"""tcgen05 GEMM vs a PyTorch fp32 reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(128, 256, 64), (256, 512, 256), (4096, 6144, 4096), (1000, 777 * 8, 520),
                                   (130, 264, 72), (8192, 4096, 14336)])
def test_gemm_bf16_nt_matches_fp32(shape):
    from deepspeed_b200.ops import kernels as K
    assert K.available()
    M, N, Kd = shape
    torch.manual_seed(0)
    a = torch.randn(M, Kd, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, Kd, device="cuda", dtype=torch.bfloat16)
    out = K.gemm_bf16_nt(a, b)
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    err = (out.float() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 1e-2 * scale + 1e-2, (err, scale)
    # also agree with cuBLAS bf16 within bf16 rounding
    cub = (a @ b.t()).float()
    assert (out.float() - cub).abs().max().item() <= 2e-2 * scale + 1e-2


@pytest.mark.parametrize("block_n", [128, 192, 256])
def test_gemm_forced_tile_widths(block_n):
    from deepspeed_b200.ops import kernels as K
    torch.manual_seed(2)
    a = torch.randn(1024, 1024, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(1000, 1024, device="cuda", dtype=torch.bfloat16)
    out = K.gemm_bf16_nt(a, b, block_n=block_n)
    ref = a.float() @ b.float().t()
    assert (out.float() - ref).abs().max().item() <= 1e-2 * ref.abs().max().item() + 1e-2


def test_gemm_strided_operands():
    from deepspeed_b200.ops import kernels as K
    torch.manual_seed(1)
    big_a = torch.randn(512, 1024, device="cuda", dtype=torch.bfloat16)
    big_b = torch.randn(768, 1024, device="cuda", dtype=torch.bfloat16)
    a, b = big_a[:, :512], big_b[:, 256:768]
    out = K.gemm_bf16_nt(a, b)
    ref = a.float() @ b.float().t()
    assert (out.float() - ref).abs().max().item() <= 1e-2 * ref.abs().max().item() + 1e-2
