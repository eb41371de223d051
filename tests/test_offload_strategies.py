"""SuperOffload (speculative CPU steps with rollback) and ZenFlow (selective GPU updates + deferred CPU update).
Mirrors reference tests/unit/runtime/zenflow and tests/unit/runtime/zero/test_zero_offloadpp.py."""
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _run(offload_extra, zero_extra=None, clip=0.0, steps=8, world_batches=True):
    import deepspeed_b200 as ds
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=32, nlayers=3)
    off = {"device": "cpu", "pin_memory": False}
    off.update(offload_extra or {})
    zc = {"stage": 2, "offload_optimizer": off, "reduce_bucket_size": 2000}
    zc.update(zero_extra or {})
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "gradient_clipping": clip,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.01}}, "zero_optimization": zc})
    losses = []
    for x, y in random_batches(steps, 4 * world, 32, "cpu", seed=1):
        loss = eng(x[rank * 4:(rank + 1) * 4], y[rank * 4:(rank + 1) * 4])
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    z = eng.optimizer
    if z._zenflow is not None:
        z._zenflow.finalize()
    return losses, z._super is not None, z._zenflow is not None


@pytest.mark.parametrize("clip", [0.0, 0.05])
def test_superoffload_matches_plain_offload(clip):
    """clip=0.05 forces the rollback path on every step (the speculative unclipped update is always wrong)."""
    a, used, _ = run_distributed(_run, 2, args=({"super_offload": True}, None, clip))[0]
    b, used_b, _ = run_distributed(_run, 2, args=(None, None, clip))[0]
    assert used and not used_b
    assert a == pytest.approx(b, abs=2e-5), (a, b)
    assert a[-1] < a[0]


def test_zenflow_full_selection_equals_adamw_and_partial_trains():
    # topk_ratio=1 with update_interval=1: every coordinate is "important" -> plain AdamW on the GPU
    full, _, used = run_distributed(_run, 1, args=(None, {"zenflow": {"topk_ratio": 1.0, "update_interval": 1,
                                                                         "select_interval": 1000}}))[0]
    ref, _, _ = run_distributed(_run, 1, args=(None, None))[0]
    assert used
    assert full == pytest.approx(ref, abs=5e-4), (full, ref)
    part, _, _ = run_distributed(_run, 2, args=(None, {"zenflow": {"topk_ratio": 0.2, "update_interval": 2,
                                                                      "select_interval": 4, "overlap_step": True}},
                                                 0.0, 16))[0]
    assert sum(part[-4:]) < sum(part[:4]), part
