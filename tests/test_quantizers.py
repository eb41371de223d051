"""Quantizer ops: CPU math paths here; the CUDA kernels are checked against the same math in test_quant_gpu.py."""
import pytest
import torch


def test_fp8_block_roundtrip_error_bound():
    from deepspeed_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = (torch.randn(8 * 512) * 3).to(torch.bfloat16)
    q, s = Q.fp8_block_quantize(x)
    y = Q.fp8_block_dequantize(q, s)
    rel = (y.float() - x.float()).abs() / (x.float().abs() + 1e-3)
    assert rel.median() < 0.04 and q.numel() == x.numel() and s.numel() == 8


@pytest.mark.parametrize("bits", [8, 4])
def test_int_group_quant_roundtrip_and_swizzle(bits):
    from deepspeed_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = torch.randn(16 * 256).to(torch.bfloat16)
    q, s = Q.quantize(x, 16, bits)
    y = Q.dequantize(q, s, 16, bits)
    tol = (x.float().abs().max() / (127 if bits == 8 else 7)).item()
    assert (y.float() - x.float()).abs().max().item() <= tol * 0.51 + 1e-2
    # swizzled layout is a permutation of the groups: [node, dev] -> [dev, node]
    qs, ss = Q.swizzle_quant(x, 16, bits, 2, 4)
    ys = Q.dequantize(qs, ss, 16, bits).view(4, 2, 2, 256).permute(1, 0, 2, 3).reshape(-1)
    torch.testing.assert_close(ys, y)


def test_quantized_reduction_means_chunks():
    from deepspeed_b200.ops import quantizer as Q
    torch.manual_seed(0)
    x = torch.randn(4 * 1024).to(torch.bfloat16)
    q, s = Q.quantize(x, 16, 8)
    out = Q.quantized_reduction(q, s, 16, 4, 8, nchunks=4)
    ref = Q.dequantize(q, s, 16, 8).float().view(4, 1024).mean(0)
    torch.testing.assert_close(out.float(), ref, rtol=1e-2, atol=1e-2)


@pytest.mark.parametrize("q_bits", [8, 6, 4, 12])
def test_fp_quantize_formats(q_bits):
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    torch.manual_seed(0)
    x = torch.randn(4, 1024).to(torch.bfloat16)
    fq = FP_Quantize(group_size=512)
    qt = fq.quantize(x, q_bits=q_bits)
    y = fq.dequantize(qt)
    err = (y.float() - x.float()).abs().max().item()
    bound = {8: 0.3, 6: 0.6, 4: 1.6, 12: 0.05}[q_bits]
    assert y.shape == x.shape and err < bound, err
    assert fq.get_scales().numel() == 8


def test_fp_selective_dequantize():
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    x = torch.randn(4, 8, 128).to(torch.bfloat16)
    fq = FP_Quantize(group_size=256)
    qt = fq.quantize(x, q_bits=8)
    sel = fq.selective_dequantize(qt, torch.tensor([2, 0]))
    full = fq.dequantize(qt)
    torch.testing.assert_close(sel, full[[2, 0]])


@pytest.mark.parametrize("eb,mb,fmax", [(3, 2, 28.0), (2, 1, 6.0), (4, 7, 510.0)])
def test_minifloat_codes_roundtrip_exhaustively(eb, mb, fmax):
    from deepspeed_b200.ops.fp_quantizer import decode_minifloat, encode_minifloat
    n = 1 << (eb + mb + 1)
    codes = torch.arange(n, dtype=torch.int32)
    vals = decode_minifloat(codes, eb, mb)
    assert vals.abs().max().item() == fmax
    back = encode_minifloat(vals, eb, mb).to(torch.int32) & (n - 1)
    assert bool(((back == codes) | (vals == 0)).all())  # -0 is stored as +0


def test_fp6_and_fp12_payloads_are_bit_packed():
    """6-bit: four e3m2 codes in three bytes, code i at bits [6i, 6i+6) (the layout csrc/wo_gemm.cu unpacks)."""
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize, pack_6bit, unpack_6bit, pack_12bit, unpack_12bit
    codes = torch.tensor([0x01, 0x3F, 0x2A, 0x15], dtype=torch.uint8)
    packed = pack_6bit(codes)
    word = 0x01 | (0x3F << 6) | (0x2A << 12) | (0x15 << 18)
    assert packed.tolist() == [word & 0xFF, (word >> 8) & 0xFF, (word >> 16) & 0xFF]
    assert unpack_6bit(packed).tolist() == codes.tolist()
    c12 = torch.tensor([0xABC, 0x123], dtype=torch.int16)
    assert unpack_12bit(pack_12bit(c12)).tolist() == c12.tolist()
    x = torch.randn(16, 256, dtype=torch.bfloat16)
    for bits, tol in ((6, 0.13), (12, 0.004)):
        fq = FP_Quantize(group_size=128)
        q = fq.quantize(x, q_bits=bits)
        assert q.payload.dtype == torch.uint8 and q.payload.numel() == x.numel() * bits // 8
        y = fq.dequantize(q)
        rel = ((y.float() - x.float()).abs().amax() / x.float().abs().amax()).item()
        assert rel < tol, (bits, rel)


def test_matmul_fp8_matches_dequantised_matmul():
    import torch
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    from deepspeed_b200.ops.fp_quantizer.fp8_gemm import matmul_fp8, matmul_fp8_fallback
    torch.manual_seed(0)
    q = FP_Quantize(group_size=64)
    w = torch.randn(128, 96, dtype=torch.bfloat16)
    wq = q.quantize(w, q_bits=8)
    scale = q.get_scales()
    x = torch.randn(4, 128, dtype=torch.bfloat16)
    want = x @ q.dequantize(wq, scale=scale)
    torch.testing.assert_close(matmul_fp8(x, wq, scale, 64, q), want)
    torch.testing.assert_close(matmul_fp8_fallback(x, wq, scale, 64, q), want)
    assert (want.float() - x.float() @ w.float()).abs().max() < 0.1 * (x.float() @ w.float()).abs().max()
