"""Activation checkpointing (runtime/activation_checkpointing.py). Mirrors the reference's
tests/unit/runtime/activation_checkpointing/test_activation_checkpointing.py: checkpointed and plain runs agree on
outputs and gradients for tensor / mixed / non-tensor arguments, dropout replays, partitioned activations at mp=2."""
import pytest
import torch

from deepspeed_b200.runtime import activation_checkpointing as ckpt
from common import run_distributed


def _reset():
    ckpt.configure(None, partition_activations=False, contiguous_checkpointing=False, checkpoint_in_cpu=False,
                   synchronize=False, profile=False)


class Masked(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(16, 16)
        self.b = torch.nn.Linear(16, 16)

    def forward(self, x, mask, scale=None):
        h = torch.tanh(self.a(x)) * mask
        if scale is not None:
            h = h * scale
        return self.b(h), mask


def _grads(mod, fn, *args):
    for p in mod.parameters():
        p.grad = None
    ins = [a.detach().clone().requires_grad_(a.requires_grad) if torch.is_tensor(a) else a for a in args]
    out = fn(*ins)
    outs = out if isinstance(out, tuple) else (out, )
    sum(o.float().sum() for o in outs if torch.is_tensor(o) and o.is_floating_point() and o.requires_grad).backward()
    return outs, [p.grad.clone() for p in mod.parameters()], [a.grad for a in ins if torch.is_tensor(a) and a.requires_grad]


def test_checkpoint_matches_plain_run_with_mixed_arguments():
    _reset()
    torch.manual_seed(0)
    m = Masked()
    x = torch.randn(4, 16, requires_grad=True)
    mask = (torch.rand(4, 16) > 0.3).float()
    for extra in ((), (2.5, )):  # a non-tensor argument goes through untouched
        o0, g0, i0 = _grads(m, m, x, mask, *extra)
        o1, g1, i1 = _grads(m, lambda *a: ckpt.checkpoint(m, *a), x, mask, *extra)
        assert torch.equal(o0[0], o1[0]) and torch.equal(o0[1], o1[1])
        assert all(torch.equal(a, b) for a, b in zip(g0, g1))
        assert all(torch.equal(a, b) for a, b in zip(i0, i1))


def test_dropout_is_replayed_in_the_recompute():
    _reset()
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(32, 32), torch.nn.Dropout(0.5), torch.nn.Linear(32, 4))
    x = torch.randn(8, 32, requires_grad=True)
    torch.manual_seed(7)
    _, g0, i0 = _grads(net, net, x)
    torch.manual_seed(7)
    _, g1, i1 = _grads(net, lambda a: ckpt.checkpoint(net, a), x)
    assert all(torch.equal(a, b) for a, b in zip(g0, g1)) and torch.equal(i0[0], i1[0])
    before = torch.get_rng_state()
    ckpt.checkpoint(net, x).sum().backward()
    assert not torch.equal(before, torch.get_rng_state())  # the outer RNG stream advances exactly once


def test_rng_tracker_forks_and_restores():
    tr = ckpt.get_cuda_rng_tracker()
    tr.reset()
    tr.add("mp", 1234)
    with pytest.raises(Exception):
        tr.add("mp", 99)  # a name can be added once
    with pytest.raises(Exception):
        tr.add("other", 1234)  # and a seed used once
    outer = torch.get_rng_state()
    with tr.fork("mp"):
        a = torch.rand(3)
    assert torch.equal(outer, torch.get_rng_state())
    saved = tr.get_states()
    tr.reset()
    tr.add("mp", 1234)
    with tr.fork("mp"):
        b = torch.rand(3)
    assert torch.equal(a, b)
    tr.set_states(saved)
    tr.reset()


def test_configure_rules_and_helpers():
    with pytest.raises(AssertionError):
        ckpt.configure(None, partition_activations=False, contiguous_checkpointing=True, num_checkpoints=2)
    _reset()
    assert ckpt.is_configured()
    ts, nts, flags = ckpt.extract_tensors((torch.ones(2), 3, "s", torch.zeros(1)))
    assert len(ts) == 2 and list(nts) == [3, "s"] and list(flags) == [True, False, False, True]
    merged = ckpt.merge_tensors(ts, nts, flags)
    assert torch.equal(merged[0], torch.ones(2)) and merged[1] == 3 and merged[2] == "s"
    d = ckpt.detach_variable((torch.ones(2, requires_grad=True), 5))
    assert d[0].requires_grad and d[0].grad_fn is None and d[1] == 5


def _partitioned():
    import deepspeed_b200.comm as dist
    rank, world = dist.get_rank(), dist.get_world_size()

    class MPU:  # every rank in one model-parallel group

        def get_model_parallel_rank(self):
            return rank

        def get_model_parallel_world_size(self):
            return world

        def get_model_parallel_group(self):
            return None

        get_data_parallel_rank = lambda self: 0
        get_data_parallel_world_size = lambda self: 1
        get_data_parallel_group = lambda self: None

    torch.manual_seed(0)
    m = Masked()
    x = torch.randn(5, 16, requires_grad=True)  # 80 elements: not a multiple of... 2 divides, use odd below
    mask = (torch.rand(5, 16) > 0.3).float()
    o0, g0, i0 = _grads(m, m, x, mask)
    ckpt.configure(MPU(), partition_activations=True)
    o1, g1, i1 = _grads(m, lambda *a: ckpt.checkpoint(m, *a), x, mask)
    assert torch.equal(o0[0], o1[0])
    assert all(torch.equal(a, b) for a, b in zip(g0, g1)) and torch.equal(i0[0], i1[0])
    odd = torch.randn(3, 7)  # 21 elements over 2 ranks: padded slice
    part = ckpt._partition(odd)
    assert part.numel() == 11
    assert torch.equal(ckpt._gather(part, odd.shape, odd.numel()), odd)
    _reset()


def test_partitioned_activations_two_ranks():
    run_distributed(_partitioned, 2)
