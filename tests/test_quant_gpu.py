"""CUDA quantization kernels vs the PyTorch reference math (same functions forced onto the CPU path)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_fp8_block_quant_kernel_matches_torch():
    from deepspeed_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = (torch.randn(64 * 512, device="cuda") * 2).to(torch.bfloat16)
    q, s = Q.fp8_block_quantize(x)
    qc, sc = Q.fp8_block_quantize(x.cpu())
    torch.testing.assert_close(s.cpu(), sc, rtol=1e-5, atol=1e-8)
    y, yc = Q.fp8_block_dequantize(q, s), Q.fp8_block_dequantize(qc, sc)
    # identical except for round-to-nearest ties resolved by the hardware converter
    assert ((y.cpu().float() - yc.float()).abs() > 1e-6).float().mean().item() < 0.02
    rel = (y.float() - x.float()).abs() / (x.float().abs() + 1e-3)
    assert rel.median().item() < 0.04


@pytest.mark.parametrize("bits", [8, 4])
def test_int_group_quant_kernels(bits):
    from deepspeed_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = torch.randn(32 * 512, device="cuda").to(torch.bfloat16)
    q, s = Q.quantize(x, 32, bits)
    qc, sc = Q.quantize(x.cpu(), 32, bits)
    torch.testing.assert_close(s.cpu(), sc, rtol=1e-5, atol=1e-8)
    y = Q.dequantize(q, s, 32, bits)
    yc = Q.dequantize(qc, sc, 32, bits)
    assert ((y.cpu().float() - yc.float()).abs() > 1e-6).float().mean().item() < 0.01
    qs, ss = Q.swizzle_quant(x, 32, bits, 2, 4)
    ys = Q.dequantize(qs, ss, 32, bits).view(4, 2, 4, 512).permute(1, 0, 2, 3).reshape(-1)
    torch.testing.assert_close(ys, y)
    out = Q.quantized_reduction(q, s, 32, 8, bits, nchunks=4)
    ref = y.float().view(4, -1).mean(0)
    torch.testing.assert_close(out.float(), ref, rtol=2e-2, atol=2e-2)


@pytest.mark.parametrize("q_bits", [8, 6, 4, 12])
def test_fp_quantizer_kernel(q_bits):
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    torch.manual_seed(0)
    x = torch.randn(16, 1024, device="cuda").to(torch.bfloat16)
    fq = FP_Quantize(group_size=512)
    y = fq.dequantize(fq.quantize(x, q_bits=q_bits))
    fqc = FP_Quantize(group_size=512)
    yc = fqc.dequantize(fqc.quantize(x.cpu(), q_bits=q_bits))
    assert ((y.cpu().float() - yc.float()).abs() > 2e-2).float().mean().item() < 0.02
