"""Asynchronous activation offload (runtime/offload_activations.py): engine bookkeeping on CPU, saved-tensor hooks,
GPU round trip. Reference: runtime/activation_checkpointing/offload_activations.py."""
import pytest
import torch

from deepspeed_b200.runtime.offload_activations import (ActivationOffloadEngine, CheckpointHiddenStatesOffload,
                                                        mark_for_offload)


def test_engine_argument_validation_and_cpu_passthrough():
    with pytest.raises(ValueError):
        ActivationOffloadEngine(keep_last_count=-1)
    with pytest.raises(ValueError):
        ActivationOffloadEngine(min_offload_bytes=-5)
    eng = ActivationOffloadEngine()
    t = torch.randn(64, 64)
    assert eng.offload(t) is t and eng.restore(t) is t  # host tensors are left alone
    assert eng.stats.saved_tensors_seen == 1 and eng.stats.offloaded_tensors == 0


def test_context_on_cpu_keeps_gradients_exact():
    torch.manual_seed(0)
    lin = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.GELU(), torch.nn.Linear(64, 8))
    x = torch.randn(16, 32, requires_grad=True)
    lin(x).square().sum().backward()
    want = [p.grad.clone() for p in lin.parameters()] + [x.grad.clone()]
    for p in lin.parameters():
        p.grad = None
    x.grad = None
    with CheckpointHiddenStatesOffload() as ctx:
        loss = lin(x).square().sum()
    loss.backward()
    got = [p.grad for p in lin.parameters()] + [x.grad]
    assert all(torch.equal(a, b) for a, b in zip(got, want))
    assert ctx.stats.saved_tensors_seen > 0 and ctx.stats.offloaded_tensors == 0


@pytest.mark.gpu
def test_gpu_offload_roundtrip_pool_and_keep_last():
    dev = torch.device("cuda")
    eng = ActivationOffloadEngine(min_offload_bytes=0, keep_last_count=1, max_in_flight=1)
    xs = [torch.randn(256, 256, device=dev) for _ in range(4)]
    tickets = [eng.offload(x) for x in xs]
    assert eng.stats.offloaded_tensors == 3 and eng.stats.kept_last_tensors == 1  # the most recent stays on the device
    small = torch.ones(4, device=dev)
    eng2 = ActivationOffloadEngine(min_offload_bytes=1024)
    assert eng2.offload(small) is small and eng2.stats.skipped_marked_tensors == 1
    outs = [eng.restore(t) for t in reversed(tickets)]
    torch.cuda.synchronize()
    for o, x in zip(outs, reversed(xs)):
        assert o.device == x.device and torch.equal(o, x)
    assert outs[0].data_ptr() == xs[3].data_ptr()  # the kept one came back without a copy
    eng.reset()
    assert eng._pool_count >= 1  # host buffers were recycled
    t2 = eng.offload(torch.randn(256, 256, device=dev))
    t3 = eng.offload(torch.randn(256, 256, device=dev))
    n_pool_after = eng._pool_count
    assert n_pool_after < 3  # a pooled pinned buffer was reused instead of allocating
    eng.restore(t3), eng.restore(t2)


@pytest.mark.gpu
def test_gpu_context_matches_plain_backward_with_checkpointing():
    from torch.utils.checkpoint import checkpoint
    dev = torch.device("cuda")
    torch.manual_seed(0)
    blocks = torch.nn.ModuleList([torch.nn.Sequential(torch.nn.Linear(256, 256), torch.nn.GELU()) for _ in range(4)]).to(dev)
    x = torch.randn(64, 256, device=dev)

    def run(offload, only_marked=False):
        for p in blocks.parameters():
            p.grad = None
        ctx = CheckpointHiddenStatesOffload(min_offload_bytes=0, only_marked=only_marked) if offload else None
        h = x
        if ctx:
            ctx.__enter__()
        for b in blocks:
            if only_marked:
                mark_for_offload(h)
            h = checkpoint(b, h, use_reentrant=False)
        loss = h.square().mean()
        if ctx:
            ctx.__exit__(None, None, None)
        loss.backward()
        torch.cuda.synchronize()
        return [p.grad.clone() for p in blocks.parameters()], (ctx.stats if ctx else None)

    want, _ = run(False)
    got, stats = run(True)
    assert all(torch.equal(a, b) for a, b in zip(got, want))
    assert stats.offloaded_tensors >= 2 and stats.restored_tensors + stats.kept_last_tensors >= 1
    got2, stats2 = run(True, only_marked=True)
    assert all(torch.equal(a, b) for a, b in zip(got2, want))
    assert 0 < stats2.marked_tensors <= 4
