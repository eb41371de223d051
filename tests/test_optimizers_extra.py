"""Muon (stand-alone and under ZeRO), 1-bit compressed all-reduce and the 1-bit optimizers (gloo, world 2).
Mirrors reference tests/unit/ops/muon, tests/unit/runtime/half_precision/onebit/test_onebit.py."""
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _muon_zero(stage):
    import deepspeed_b200 as ds
    import torch.distributed as dist
    from deepspeed_b200.ops.muon import Muon
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=16, nlayers=2)
    ref = SimpleModel(hidden_dim=16, nlayers=2)
    ref.load_state_dict(model.state_dict())
    params = {"lr": 0.02, "momentum": 0.9, "weight_decay": 0.01, "adam_lr": 1e-2, "betas": [0.9, 0.95], "eps": 1e-8,
              "ns_method": "standard"}
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "Muon", "params": params},
        "zero_optimization": {"stage": stage}})
    ropt = Muon(ref.parameters(), **params)
    losses, rlosses = [], []
    for x, y in random_batches(5, 4 * world, 16, "cpu", seed=3):
        loss = eng(x[rank * 4:(rank + 1) * 4], y[rank * 4:(rank + 1) * 4])
        eng.backward(loss)
        eng.step()
        rl = ref(x, y)
        ropt.zero_grad()
        rl.backward()
        ropt.step()
        t = loss.detach().clone()
        dist.all_reduce(t)
        losses.append(float(t) / world)
        rlosses.append(float(rl))
    return losses, rlosses


@pytest.mark.parametrize("stage", [1, 3])
def test_muon_under_zero_matches_single_process(stage):
    outs = run_distributed(_muon_zero, 2, args=(stage, ))
    losses, ref = outs[0]
    for a, b in zip(losses, ref):
        assert abs(a - b) < 2e-3, (losses, ref)
    assert losses[-1] < losses[0]


def _compressed():
    import torch.distributed as dist
    from deepspeed_b200.runtime.comm.compressed import CompressedBackend, pack_signs, unpack_signs
    rank, world = dist.get_rank(), dist.get_world_size()
    x = torch.randn(64)
    assert torch.equal(unpack_signs(pack_signs(x)), torch.where(x >= 0, 1.0, -1.0))
    be = CompressedBackend()
    n = 1000
    we = torch.zeros(be.padded_size(n))
    se = torch.zeros(be.padded_size(n) // world)
    torch.manual_seed(rank)
    base = torch.randn(n)
    # error feedback: the running sum of the compressed results tracks the running sum of the exact means
    acc_c, acc_e = torch.zeros(n), torch.zeros(n)
    for it in range(200):
        t = base + 0.1 * torch.randn(n)
        exact = t.clone()
        dist.all_reduce(exact)
        exact /= world
        acc_e += exact
        acc_c += be.compressed_allreduce(t, we, se)
    rel = float((acc_c - acc_e).norm() / acc_e.norm())
    same = acc_c.clone()
    dist.broadcast(same, 0)
    return rel, bool(torch.equal(same, acc_c))


def test_compressed_allreduce_error_feedback():
    outs = run_distributed(_compressed, 2)
    for rel, same in outs:
        assert rel < 0.05, rel  # 1 bit per element, yet the accumulated error stays bounded
        assert same             # every rank reconstructs the identical tensor


def _onebit(name):
    import deepspeed_b200 as ds
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=16, nlayers=2)
    params = {"lr": 1e-3 if name == "ZeroOneAdam" else 1e-2, "weight_decay": 0.0}
    params.update({"OneBitAdam": {"freeze_step": 4}, "OneBitLamb": {"freeze_step": 4},
                   "ZeroOneAdam": {"var_freeze_step": 12, "var_update_scaler": 2, "local_step_scaler": 2,
                                   "local_step_clipper": 2}}[name])
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "optimizer": {"type": name, "params": params},
        "zero_optimization": {"stage": 0}})
    losses = []
    data = random_batches(4, 4 * world, 16, "cpu", seed=3)
    for step in range(24):
        x, y = data[step % 4]
        loss = eng(x[rank * 4:(rank + 1) * 4], y[rank * 4:(rank + 1) * 4])
        eng.backward(loss)
        eng.step()
        losses.append(float(loss))
    inner = eng.optimizer.optimizer
    w = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    w0 = w.clone()
    dist.broadcast(w0, 0)
    return losses, inner.compression_active, float((w - w0).abs().max())


@pytest.mark.parametrize("name", ["OneBitAdam", "ZeroOneAdam", "OneBitLamb"])
def test_onebit_optimizers_train_and_stay_in_sync(name):
    outs = run_distributed(_onebit, 2, args=(name, ))
    for losses, active, drift in outs:
        assert active, "compression stage was never entered"
        assert sum(losses[-4:]) < sum(losses[:4]), losses
        assert drift < 1e-5, drift  # replicas stay identical although gradients are no longer all-reduced
