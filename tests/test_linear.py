"""deepspeed.linear: OptimizedLinear factory, LoRA adapters, quantized frozen weights, base-weight sharding.
Mirrors reference tests/unit/linear/{test_linear.py,test_quant_param.py,test_ctx.py}."""
import copy

import pytest
import torch

from common import run_distributed
from deepspeed_b200.linear import (Init, LoRAConfig, LoRAOptimizedLinear, OptimizedLinear, QuantizationConfig,
                                   QuantizedLinear, QuantizedParameter)


def test_factory_and_lora_identity_at_init():
    torch.manual_seed(0)
    assert type(OptimizedLinear(8, 8)) is torch.nn.Linear
    layer = OptimizedLinear(64, 32, lora_config=LoRAConfig(lora_r=4, lora_alpha=8), dtype=torch.float32)
    assert isinstance(layer, LoRAOptimizedLinear) and not layer.weight.requires_grad
    x = torch.randn(5, 64)
    torch.testing.assert_close(layer(x), torch.nn.functional.linear(x, layer.weight))  # B = 0 at init
    layer.lora_weight_2.weight.data.normal_()
    want = torch.nn.functional.linear(x, layer.weight) + 2.0 * layer.lora_weight_2(layer.lora_weight_1(x))
    torch.testing.assert_close(layer(x), want)
    layer(x).sum().backward()
    assert layer.weight.grad is None and layer.lora_weight_1.weight.grad is not None


@pytest.mark.parametrize("q_bits,tol", [(8, 0.07), (6, 0.13), (4, 0.3)])
def test_quantized_parameter_roundtrip(q_bits, tol):
    torch.manual_seed(1)
    w = torch.randn(64, 128, dtype=torch.bfloat16)
    qp = QuantizedParameter(w, quantization_config=QuantizationConfig(q_bits=q_bits, group_size=128))
    assert qp.dtype in (torch.uint8, torch.bfloat16) and not qp.requires_grad
    if q_bits == 8:
        assert qp.numel() == w.numel() and qp.dtype == torch.uint8
    if q_bits == 4:
        assert qp.numel() == w.numel() // 2
    d = qp.dequantized()
    assert d.shape == w.shape and d.dtype == torch.bfloat16
    rel = ((d.float() - w.float()).abs().max() / w.float().abs().max()).item()
    assert rel < tol, rel
    c = copy.deepcopy(qp)
    assert torch.equal(c.dequantized(), d)
    with pytest.raises(ValueError):
        QuantizedParameter(w, requires_grad=True)


def test_quantized_linear_and_lora_with_quant():
    torch.manual_seed(2)
    ql = QuantizedLinear(128, 64, quantization_config=QuantizationConfig(q_bits=8, group_size=128))
    x = torch.randn(3, 128, dtype=torch.bfloat16)
    ref = torch.nn.functional.linear(x, ql.weight.dequantized())
    torch.testing.assert_close(ql(x), ref)
    lq = OptimizedLinear(128, 64, lora_config=LoRAConfig(lora_r=4), quantization_config=QuantizationConfig(group_size=128))
    assert isinstance(lq.weight, QuantizedParameter)
    lq(x).float().sum().backward()
    assert lq.lora_weight_2.weight.grad is not None


def test_init_context_swaps_linear():
    with Init(lora_config=LoRAConfig(lora_r=2, target_mods=["proj"]), quant_config=None):
        lin = torch.nn.Linear(16, 8, bias=False, dtype=torch.float32)
    assert isinstance(lin, LoRAOptimizedLinear) and not lin._initialized  # delayed until init_lora(model)
    assert torch.nn.Linear is not type(lin) and isinstance(torch.nn.Linear(4, 4), torch.nn.Linear)
    from deepspeed_b200.linear import init_lora
    m = torch.nn.Sequential(lin)
    init_lora(m)
    assert lin._initialized and lin.lora_weight_1.weight.requires_grad and not lin.weight.requires_grad


def _sharded():
    import torch.distributed as dist
    torch.manual_seed(0)
    world, rank = dist.get_world_size(), dist.get_rank()
    full = torch.randn(32, 64)
    layer = LoRAOptimizedLinear(64, 32, lora_config=LoRAConfig(lora_r=4, base_weight_sharding=world,
                                                               target_mods=["proj"]), dtype=torch.float32, device="cpu")
    assert layer.weight.numel() == 32 * 64 // world
    sd = {"proj.weight": full.clone(), "proj.lora_weight_1.weight": layer.lora_weight_1.weight.data.clone(),
          "proj.lora_weight_2.weight": layer.lora_weight_2.weight.data.clone()}
    holder = torch.nn.ModuleDict({"proj": layer})
    holder.load_state_dict(sd)
    x = torch.randn(3, 64)
    return float((layer(x) - torch.nn.functional.linear(x, full)).abs().max())


def test_base_weight_sharding_gathers_full_weight():
    for err in run_distributed(_sharded, 2):
        assert err < 1e-5
