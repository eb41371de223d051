"""Fused NVLink MoE dispatch / combine (csrc/moe_comm.cu) vs the NCCL all-to-all-v path of the same layer, 2 GPUs."""
import os

import pytest
import torch

from common import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")


def _run_layer(fused, top_k, n_experts):
    import deepspeed_b200 as ds
    from deepspeed_b200.parallel.auto_ep import AutoEPMoELayer
    os.environ["DSB200_FUSED_MOE"] = "1" if fused else "0"
    dev = torch.device("cuda", torch.cuda.current_device())
    rank = ds.comm.get_rank()
    torch.manual_seed(0)
    H, I = 256, 512
    layer = AutoEPMoELayer(H, I, n_experts, top_k, ep_size=2, dtype=torch.bfloat16, device=dev)
    layer.router.to(dev)
    torch.manual_seed(100 + rank)  # every rank routes different tokens, different counts per expert
    T = 300 + 17 * rank
    x = torch.randn(1, T, H, device=dev, dtype=torch.bfloat16, requires_grad=True)
    outs = []
    for it in range(2):  # two calls: staging buffers are reused (epoch protocol)
        y = layer(x)
        (y.float() * torch.linspace(-1, 1, H, device=dev)).sum().backward()
        outs.append(y.detach().float().clone())
    used = getattr(layer, "_fused_ok", False)
    return (outs[-1].cpu(), x.grad.float().cpu(), layer.experts.w13.grad.float().cpu(), layer.experts.w2.grad.float().cpu(),
            layer.router.weight.grad.float().cpu() if hasattr(layer.router, "weight") else None, used)


@needs2
@pytest.mark.parametrize("top_k,n_experts", [(2, 4), (1, 2), (2, 8)])
def test_fused_moe_matches_nccl_path(top_k, n_experts):
    a = run_distributed(_run_layer, 2, args=(True, top_k, n_experts))
    b = run_distributed(_run_layer, 2, args=(False, top_k, n_experts))
    for ra, rb in zip(a, b):
        assert ra[5] is True and rb[5] is False
        for i, (ta, tb) in enumerate(zip(ra[:5], rb[:5])):
            if ta is None:
                continue
            scale = max(1.0, tb.abs().max().item())
            assert (ta - tb).abs().max().item() <= 3e-2 * scale, (i, (ta - tb).abs().max().item(), scale)
