"""Pipeline: partitioning, 1F1B training on 2 stages (CPU/gloo) equals the single-stage result
(reference tests/unit/runtime/pipe/test_pipe.py, test_pipe_schedule.py, test_topology.py)."""
import torch
import torch.nn as nn

from common import run_distributed


def test_partition_helpers():
    from deepspeed_b200.runtime.pipe.module import partition_balanced, partition_uniform
    assert partition_uniform(10, 4) == [0, 3, 6, 8, 10]
    p = partition_balanced([1, 1, 1, 10, 1, 1], 2)
    assert p[0] == 0 and p[-1] == 6 and max(sum([1, 1, 1, 10, 1, 1][p[i]:p[i + 1]]) for i in range(2)) <= 13


def test_train_schedule_invariants():
    from deepspeed_b200.runtime.pipe import schedule as S
    for M, stages in ((4, 2), (8, 4), (3, 3)):
        for s in range(stages):
            cmds = [c for step in S.TrainSchedule(M, stages, s).steps() for c in step]
            assert sum(isinstance(c, S.ForwardPass) for c in cmds) == M
            assert sum(isinstance(c, S.BackwardPass) for c in cmds) == M
            assert isinstance(cmds[-1], S.OptimizerStep)
            # a buffer is never forwarded twice before its backward
            live = set()
            for c in cmds:
                if isinstance(c, S.ForwardPass):
                    assert c.buffer_id not in live
                    live.add(c.buffer_id)
                if isinstance(c, S.BackwardPass):
                    live.discard(c.buffer_id)


def _layers():
    torch.manual_seed(0)
    return [nn.Linear(8, 16), nn.Tanh(), nn.Linear(16, 16), nn.Tanh(), nn.Linear(16, 4)]


def _pipe_train(stages):
    import deepspeed_b200 as ds
    from deepspeed_b200.runtime.pipe import PipelineModule
    layers = _layers()
    model = PipelineModule(layers, num_stages=stages, loss_fn=nn.CrossEntropyLoss(), partition_method="uniform")
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": 4,
        "optimizer": {"type": "SGD", "params": {"lr": 0.1}}, "zero_optimization": {"stage": 0}},
        model_parameters=[p for p in model.parameters()])
    g = torch.Generator().manual_seed(1)
    data = [(torch.randn(4, 8, generator=g), torch.randint(0, 4, (4, ), generator=g)) for _ in range(8)]
    losses = []
    it = iter(data)
    for _ in range(2):
        losses.append(float(eng.train_batch(it)))
    first = {n: p.detach().clone() for n, p in model.named_parameters()}
    return losses, first, eng.stage_id


def test_pipeline_two_stages_matches_one():
    one = run_distributed(_pipe_train, 1, args=(1, ))[0]
    two = run_distributed(_pipe_train, 2, args=(2, ))
    for a, b in zip(one[0], two[0][0]):
        assert abs(a - b) < 1e-5, (one[0], two[0][0])
    merged = {}
    for losses, params, stage in two:
        merged.update(params)
    for n, p in one[1].items():
        torch.testing.assert_close(merged[n], p, rtol=1e-5, atol=1e-6)


def _p2p_objects():
    import deepspeed_b200.comm as dist
    from deepspeed_b200.runtime.pipe import p2p
    msg = {"step": 3, "names": ["a", "b"], "lr": 0.5, "nested": {"k": [1, 2, 3]}}
    if dist.get_rank() == 0:
        p2p.send_obj(msg, 1)
        return True
    got = p2p.recv_obj(0)
    p2p.wait()
    return got == msg and p2p.can_send_recv()


def test_p2p_send_and_receive_python_objects():
    from common import run_distributed
    assert all(run_distributed(_p2p_objects, 2))
