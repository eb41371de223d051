"""Weight-only quantized decode GEMM (csrc/wo_gemm.cu) vs dequantize + fp32 matmul."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("q_bits", [8, 4])
@pytest.mark.parametrize("M,N,K,group", [(1, 512, 1024, 256), (4, 1000, 4096, 512), (8, 96, 14336, 512), (3, 64, 256, 256)])
def test_wo_gemv_matches_dequant_matmul(q_bits, M, N, K, group):
    from deepspeed_b200.ops import kernels as Kn
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    torch.manual_seed(0)
    dev = torch.device("cuda")
    w = torch.randn(N, K, device=dev, dtype=torch.bfloat16) * 0.05
    x = torch.randn(M, K, device=dev, dtype=torch.bfloat16)
    fq = FP_Quantize(group_size=group)
    q = fq.quantize(w, q_bits=q_bits)
    w_deq = fq.dequantize(q).float()
    want = x.float() @ w_deq.t()
    payload = q.payload.view(torch.uint8)
    assert Kn.wo_gemv_supported(x, payload, group, q_bits, K)
    got = Kn.wo_gemv(x, payload, q.scale, N, K, group, q_bits)
    torch.cuda.synchronize()
    err = (got.float() - want).abs().max().item()
    assert err <= 2e-2 * max(1.0, want.abs().max().item()), err


@pytest.mark.parametrize("q_bits", [8, 6, 4])
@pytest.mark.parametrize("T,N,K,group", [(1, 512, 1024, 256), (5, 1000, 4096, 512), (16, 200, 14336, 512),
                                         (33, 4096, 4096, 128), (64, 640, 2048, 64), (7, 384, 1536, 1536)])
def test_wo_gemm_tcgen05_matches_dequant_matmul(q_bits, T, N, K, group):
    from deepspeed_b200.ops import kernels as Kn
    from deepspeed_b200.ops.fp_quantizer import FP_Quantize
    torch.manual_seed(0)
    dev = torch.device("cuda")
    w = torch.randn(N, K, device=dev, dtype=torch.bfloat16) * 0.05
    x = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
    fq = FP_Quantize(group_size=group)
    q = fq.quantize(w, q_bits=q_bits)
    want = x.float() @ fq.dequantize(q).float().t()
    payload = q.payload.view(torch.uint8)
    assert Kn.wo_gemm_supported(x, payload, group, q_bits, K)
    bias = torch.randn(N, device=dev, dtype=torch.bfloat16)
    got = Kn.wo_gemm(x, payload, q.scale, N, K, group, q_bits, bias)
    torch.cuda.synchronize()
    err = (got.float() - (want + bias.float())).abs().max().item()
    assert err <= 2e-2 * max(1.0, want.abs().max().item()), err


def test_quantized_linear_decode_uses_wo_gemv():
    from deepspeed_b200.linear import QuantizedLinear
    from deepspeed_b200.linear.config import QuantizationConfig
    from deepspeed_b200.ops import kernels as Kn
    torch.manual_seed(1)
    dev = torch.device("cuda")
    lin = QuantizedLinear(1024, 768, quantization_config=QuantizationConfig(q_bits=8, group_size=512)).to(dev)
    x = torch.randn(2, 1, 1024, device=dev, dtype=torch.bfloat16)
    with torch.no_grad():
        n0 = Kn.launch_count()
        y = lin(x)
        assert Kn.launch_count() > n0
        ref = torch.nn.functional.linear(x, lin.weight.dequantized())
    assert (y.float() - ref.float()).abs().max().item() < 3e-2 * max(1.0, ref.float().abs().max().item())
