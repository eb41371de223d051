"""ZeRO stage 0-3 parity against plain PyTorch on 2 ranks (CPU/gloo), incl. GAS, clipping, fp32/bf16
(reference test model: ``tests/unit/v1/zero/test_zero.py``)."""
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches

HID = 16


def _train(stage, dtype_name, gas, steps, world):
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = SimpleModel(HID, 3)
    cfg = {
        "train_micro_batch_size_per_gpu": 4,
        "gradient_accumulation_steps": gas,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.01}},
        "zero_optimization": {"stage": stage, "unit_max_numel": 300, "stage3_param_persistence_threshold": 20,
                              "reduce_bucket_size": 400},
        "gradient_clipping": 1.0,
    }
    if dtype_name == "bf16":
        cfg["bf16"] = {"enabled": True}
    eng, _, _, _ = ds.initialize(model=model, config=cfg)
    rank = ds.comm.get_rank()
    dt = torch.bfloat16 if dtype_name == "bf16" else torch.float32
    losses = []
    batches = random_batches(steps * gas * world, 4, HID, "cpu", dtype=dt)
    for i in range(steps * gas):
        x, y = batches[i * world + rank]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param
    full = {n: safe_get_full_fp32_param(p).clone() for n, p in eng.module.named_parameters()}
    return losses, full, eng.get_global_grad_norm()


def _reference(gas, steps, world, dtype_name):
    torch.manual_seed(0)
    model = SimpleModel(HID, 3)
    opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.01)
    batches = random_batches(steps * gas * world, 4, HID, "cpu")
    for s in range(steps):
        opt.zero_grad()
        for i in range(s * gas * world, (s + 1) * gas * world):
            x, y = batches[i]
            (model(x, y) / (gas * world)).backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
    return {n: p.detach().clone() for n, p in model.named_parameters()}


@pytest.mark.parametrize("stage", [0, 1, 2, 3])
@pytest.mark.parametrize("gas", [1, 2])
def test_zero_fp32_matches_torch(stage, gas):
    world, steps = 2, 3
    outs = run_distributed(_train, world, args=(stage, "fp32", gas, steps, world))
    ref = _reference(gas, steps, world, "fp32")
    for rank_out in outs:
        _, full, _ = rank_out
        for n, p in ref.items():
            torch.testing.assert_close(full[n], p, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("stage", [1, 3])
def test_zero_bf16_tracks_fp32(stage):
    world, steps = 2, 3
    outs = run_distributed(_train, world, args=(stage, "bf16", 1, steps, world))
    ref = _reference(1, steps, world, "bf16")
    _, full, _ = outs[0]
    for n, p in ref.items():
        torch.testing.assert_close(full[n], p, rtol=5e-2, atol=5e-2)
    # both ranks agree exactly on the master weights
    for n in full:
        torch.testing.assert_close(outs[0][1][n], outs[1][1][n], rtol=0, atol=0)


def _llama_direct_wgrad(stage, direct):
    """Tiny Llama under ZeRO (gloo, fp32): weight gradients written straight into the flat gradient buffers
    (ops.fused.Linear + ZeroShardedOptimizer.grad_target) vs the grad-hook copy path."""
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=128, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=64, dtype=torch.float32)
    model = LlamaForCausalLM(cfg)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2, "gradient_clipping": 1.0,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
        "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 10, "unit_max_numel": 6000,
                              "reduce_bucket_size": 5000}})
    eng.optimizer._direct_wgrad = direct
    calls = {"n": 0}
    orig = eng.optimizer.grad_written

    def counted(p):
        calls["n"] += 1
        return orig(p)

    eng.optimizer.grad_written = counted
    torch.manual_seed(1 + ds.comm.get_rank())
    losses = []
    for _ in range(6):
        ids = torch.randint(0, 128, (2, 16))
        loss = eng(ids, labels=ids)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    return losses, calls["n"]


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_direct_weight_gradients_match_hook_path(stage):
    a = run_distributed(_llama_direct_wgrad, 2, args=(stage, True))
    b = run_distributed(_llama_direct_wgrad, 2, args=(stage, False))
    (la, na), (lb, nb) = a[0], b[0]
    assert na > 0 and nb == 0, (na, nb)   # the direct path really ran (and only when enabled)
    for x, y in zip(la, lb):
        assert abs(x - y) < 1e-5, (la, lb)


def _coordinator_knobs(reuse, bucket):
    """ZeRO-3 coordinator knobs: stage3_max_reuse_distance keeps the last units gathered from forward into backward,
    stage3_prefetch_bucket_size bounds how far ahead gathers are launched."""
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = SimpleModel(16, 6)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
        "zero_optimization": {"stage": 3, "unit_max_numel": 300, "stage3_param_persistence_threshold": 0,
                              "stage3_max_reuse_distance": reuse, "stage3_prefetch_bucket_size": bucket,
                              "stage3_max_live_parameters": 10**9}})
    co = eng.optimizer.coordinator
    batches = random_batches(8, 4, 16, "cpu")
    losses, per_step = [], []
    for i in range(4):
        g0 = co.gathers_launched
        x, y = batches[2 * i + ds.comm.get_rank()]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
        per_step.append(co.gathers_launched - g0)
    n_units = len([u for u in eng.optimizer.units if not u.persistent])
    return losses, per_step, n_units


def test_zero3_reuse_distance_and_prefetch_bucket():
    base = run_distributed(_coordinator_knobs, 2, args=(0, 1))[0]
    keep = run_distributed(_coordinator_knobs, 2, args=(10**9, 10**9))[0]
    some = run_distributed(_coordinator_knobs, 2, args=(700, 1))[0]
    n = base[2]
    assert base[1][-1] == 2 * n                       # every unit gathered for forward and again for backward
    assert keep[1][-1] == n                           # everything within reuse distance: one gather per step
    assert n < some[1][-1] < 2 * n                    # only the last units stay gathered
    for a, b, c in zip(base[0], keep[0], some[0]):    # scheduling does not change the math
        assert abs(a - b) < 1e-6 and abs(a - c) < 1e-6


def _sanity_and_profiler():
    import deepspeed_b200 as ds
    from deepspeed_b200.runtime.zero import partitioned_param_profiler as pp
    torch.manual_seed(0)
    model = SimpleModel(16, 4)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
        "zero_optimization": {"stage": 3, "unit_max_numel": 300, "stage3_param_persistence_threshold": 0,
                              "stage3_max_reuse_distance": 0, "enable_sanity_checks": True,
                              "log_trace_cache_warnings": True}})
    co = eng.optimizer.coordinator
    assert co.sanity_checks and co.param_profiler is not None
    batches = random_batches(8, 4, 16, "cpu")
    for i in range(3):
        loss = eng(*batches[2 * i + ds.comm.get_rank()])
        eng.backward(loss)
        eng.step()
    ev = co.param_profiler.event_counters
    n = len([u for u in eng.optimizer.units if not u.persistent])
    counts = {k: v.count for k, v in ev.items()}
    # a rank that gathers a different unit sequence is caught at the next step boundary
    caught = False
    if ds.comm.get_rank() == 1:
        co._gather_log.append(12345)
    try:
        co.check_gather_order()
    except RuntimeError as e:
        caught = "disagreement" in str(e)
    co.param_profiler.log_events()
    return counts, n, caught


def test_zero3_sanity_checks_and_param_profiler():
    res = run_distributed(_sanity_and_profiler, 2)
    counts, n, _ = res[0]
    from deepspeed_b200.runtime.zero import partitioned_param_profiler as pp
    assert counts[pp.FORWARD_ALL_GATHER] == 3 * n and counts[pp.BACKWARD_ALL_GATHER] == 3 * n
    assert counts[pp.FORWARD_FETCH_WAIT] >= 3 * n
    assert counts.get(pp.FORWARD_PREFETCH_SUBMIT, 0) > 0  # steps after the trace step prefetch one unit ahead
    assert res[1][2] and res[0][2]  # the disagreement is raised on every rank (no rank is left in a collective)


def test_mics_rank_tables():
    from deepspeed_b200.runtime.zero.mics import mics_rank_tables
    t = mics_rank_tables(8, 4)
    assert t["shard_groups"] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert t["replicate_groups"] == [[0, 4], [1, 5], [2, 6], [3, 7]] and t["span_nodes"] == 1
    t = mics_rank_tables(16, 8, ndev_per_node=4)  # a shard group over two nodes of four devices
    assert t["span_nodes"] == 2
    assert t["intra_node_groups"][:2] == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert t["inter_node_groups"][:4] == [[0, 4], [1, 5], [2, 6], [3, 7]]
    import pytest
    with pytest.raises(AssertionError):
        mics_rank_tables(8, 3)


def _mics_hier():
    import torch
    import deepspeed_b200.comm as dist
    from deepspeed_b200.runtime.zero.mics import (MiCS_AllGatherCoalescedHandle, create_mics_comm_groups,
                                                  hierarchical_all_gather)
    rank = dist.get_rank()
    g = create_mics_comm_groups(4, hierarchical_allgather=True, ndev_per_node=2)  # 4 ranks = 2 "nodes" x 2 devices
    assert (g.param_shard_size, g.param_shard_rank, g.param_repli_size) == (4, rank, 1)
    assert g.param_intra_node_group is not None and g.param_inter_node_shard_group is not None
    out = torch.empty(4 * 3)
    hierarchical_all_gather(out, torch.full((3, ), float(rank)), g)
    assert out.tolist() == [0.0] * 3 + [1.0] * 3 + [2.0] * 3 + [3.0] * 3
    g2 = create_mics_comm_groups(2)  # shard pairs, replicate across them
    assert (g2.param_shard_size, g2.param_repli_size, g2.param_repli_rank) == (2, 2, rank // 2)
    out2 = torch.empty(2)
    hierarchical_all_gather(out2, torch.tensor([float(rank)]), g2)
    assert out2.tolist() == ([0.0, 1.0] if rank < 2 else [2.0, 3.0])
    h = MiCS_AllGatherCoalescedHandle(None, [])
    h.wait()
    return h.complete


def test_mics_groups_and_hierarchical_gather_four_ranks():
    assert all(run_distributed(_mics_hier, 4))


def test_reference_zero_class_names():
    import torch
    from deepspeed_b200.runtime.zero.mics import DeepSpeedZeroOptimizer, DeepSpeedZeroOptimizer_Stage3, MiCS_Optimizer
    from deepspeed_b200.runtime.zero import ZeroShardedOptimizer
    m = torch.nn.Linear(8, 8)
    z = DeepSpeedZeroOptimizer(m, torch.optim.AdamW(m.parameters(), lr=1e-3), partition_gradients=True,
                               dtype=torch.float32)
    assert isinstance(z, ZeroShardedOptimizer) and z.stage == 2
    assert issubclass(MiCS_Optimizer, DeepSpeedZeroOptimizer_Stage3)


def test_gather_handles_and_registry():
    from deepspeed_b200.runtime.zero import (AllGatherCoalescedHandle, AllGatherHandle, InflightParamRegistry,
                                             MultipleAllGatherHandles, NoGatherCoalescedHandle, NoGatherHandle,
                                             PartitionedParameterCoordinator, ZeroParamStatus, ZeRoTraceMode,
                                             ZeROOrderedDict)

    class W:
        waited = 0

        def wait(self):
            W.waited += 1

    def param(numel=6, per=3):
        p = torch.nn.Parameter(torch.empty(0))
        p.ds_status, p.ds_numel, p.ds_shape, p.ds_id = ZeroParamStatus.INFLIGHT, numel, (2, 3), 0
        p.ds_tensor = torch.empty(per)
        p.ds_tensor.ds_numel = per
        return p

    p = param()
    AllGatherHandle(W(), p).wait()
    assert p.ds_status == ZeroParamStatus.AVAILABLE and W.waited == 1
    with pytest.raises(RuntimeError):
        AllGatherHandle(W(), p)  # not inflight any more
    a, b = param(), param(4, 2)
    b.ds_shape = (4, )
    parts = [torch.tensor([0., 1, 2, 10, 11]), torch.tensor([3., 4, 5, 12, 13])]  # rank-major partitions of a then b
    h = AllGatherCoalescedHandle(W(), [a, b], parts, 2)
    MultipleAllGatherHandles([h]).wait()
    assert a.data.tolist() == [[0, 1, 2], [3, 4, 5]] and b.data.tolist() == [10, 11, 12, 13] and h.complete
    c = param()
    NoGatherHandle(c).wait()
    d = param()
    NoGatherCoalescedHandle([d]).wait()
    assert c.ds_status == d.ds_status == ZeroParamStatus.AVAILABLE
    reg = InflightParamRegistry()
    e = param()
    reg[e] = h
    with pytest.raises(RuntimeError):
        reg[e] = h
    with pytest.raises(RuntimeError):
        reg[c] = h  # available, not inflight
    assert ZeRoTraceMode.RECORD.value == 1 and PartitionedParameterCoordinator.__name__ == "UnitCoordinator"
    mod = torch.nn.Linear(2, 2)
    od = ZeROOrderedDict(mod, mod._parameters)
    gathered = []
    w = od["weight"]
    w.ds_status = ZeroParamStatus.NOT_AVAILABLE
    w.all_gather = lambda: gathered.append(1)
    od["weight"]
    assert not gathered  # outside forward nothing is fetched
    od._in_forward = True
    od["weight"]
    assert gathered == [1]


def test_zenflow_and_superoffload_class_fronts():
    from deepspeed_b200.ops.adam.multi_tensor_apply import MultiTensorApply
    from deepspeed_b200.runtime.superoffload.superoffload_stage3 import SuperOffloadOptimizer_Stage3
    from deepspeed_b200.runtime.superoffload.superoffload_utils import EventTypes, ResultKeys, TaskKeys
    from deepspeed_b200.runtime.zenflow.zenflow_stage_1_and_2 import (ZenFlowZeroOptimizer,
                                                                      ZenFlowZeroOptimizerParallel,
                                                                      ZenFlowZeroOptimizerSequential)

    class Z:
        overlap_step = True

    assert ZenFlowZeroOptimizer.create(Z()) is ZenFlowZeroOptimizerParallel
    Z.overlap_step = False
    assert ZenFlowZeroOptimizer.create(Z()) is ZenFlowZeroOptimizerSequential
    m = torch.nn.Linear(4, 4)
    with pytest.raises(AssertionError):
        ZenFlowZeroOptimizer(m, None, zenflow=None, offload_optimizer=None)
    with pytest.raises(AssertionError):
        SuperOffloadOptimizer_Stage3(m, None, offload_optimizer=None)
    assert (TaskKeys.PARAM_GRAD, ResultKeys.UPDATED_PARAM, EventTypes.ROLLBACK) == ("param_grad", "updated_param",
                                                                                     "rollback")
    got = MultiTensorApply(2048)(lambda chunk, flag, lists, lr: (chunk, len(lists), lr), None, [[1], [2]], 0.1)
    assert got == (2048, 2, 0.1)


def test_memory_estimators(capsys):
    """Values of the reference's documented examples (docs/code-docs/source/memory.rst)."""
    from deepspeed_b200.runtime.zero import (estimate_zero2_model_states_mem_needs,
                                             estimate_zero2_model_states_mem_needs_all_live,
                                             estimate_zero3_model_states_mem_needs,
                                             estimate_zero3_model_states_mem_needs_all_cold)
    P = 2_851_000_000
    cpu, gpu = estimate_zero2_model_states_mem_needs(P, 8, 1, cpu_offload=True)
    assert gpu == 2 * P and cpu == int(P * 32 * 1.5)
    cpu, gpu = estimate_zero2_model_states_mem_needs(P, 8, 1, cpu_offload=False)
    assert gpu == 2 * P + int(18 * P / 8) and cpu == int(P * 4 * 8 * 1.5)
    L = 32_000_000
    cpu, gpu, big = estimate_zero3_model_states_mem_needs(P, L, 8, 1, True, True, True)
    assert (gpu, big) == (4 * L, 4 * L) and cpu == int(P * 18 * 1.5)
    cpu, gpu, _ = estimate_zero3_model_states_mem_needs(P, L, 8, 2, True, False, False)
    assert gpu == 4 * L + int(2 * P / 16) and cpu == int(P * max(32, 8.0) * 1.5)
    cpu, gpu, _ = estimate_zero3_model_states_mem_needs(P, L, 8, 1, False, False, True)
    assert gpu == 4 * L + int(18 * P / 8) and cpu == int(L * 32 * 1.5)
    rows = estimate_zero3_model_states_mem_needs_all_cold(P, L, 8, 1)
    assert len(rows) == 6 and "offload_param=cpu , offload_optimizer=cpu , zero_init=1" in rows[0][2]
    m = torch.nn.Sequential(torch.nn.Linear(10, 10), torch.nn.Linear(10, 10))
    m[1].weight = m[0].weight  # shared: counted once
    rows = estimate_zero2_model_states_mem_needs_all_live(m, 1, 1)
    assert rows[0][1] == 2 * 120
    out = capsys.readouterr().out
    assert "per CPU  |  per GPU" in out and "2851M total params, 32M largest layer params" in out
    from deepspeed_b200.runtime.zero.sharded_optimizer import estimate_zero3_model_states_mem_needs_all_live  # noqa: F401
