"""DeepCompile counterpart: native unit-timeline planner (csrc/cpu/dc_planner.cpp), the unit profiler and the
plan-driven ZeRO-3 coordinator. Reference behaviour: compile/passes/prefetch.py, selective_gather.py."""
import numpy as np
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _timeline(n, unit_mb=100.0, step_ms=1.0, peak_gb=10.0):
    unit_bytes = [int(unit_mb * 2**20)] * n
    step_unit = list(range(n)) + list(range(n - 1, -1, -1))
    step_us = [step_ms * 1e3] * (2 * n)
    step_peak = [int(peak_gb * 2**30)] * (2 * n)
    return unit_bytes, step_unit, step_us, step_peak


def test_fit_comm_recovers_latency_and_bandwidth():
    from deepspeed_b200.compile import fit_comm
    nbytes = [1 << 20, 4 << 20, 16 << 20, 64 << 20, 256 << 20]
    us = [15.0 + b / 250e3 for b in nbytes]
    lat, bpu = fit_comm(nbytes, us)
    assert abs(lat - 15.0) < 1e-3 and abs(bpu - 250e3) / 250e3 < 1e-6
    lat, bpu = fit_comm([1 << 20] * 3, [10.0, 11.0, 12.0])  # one size only: bandwidth-only model
    assert lat == 0 and abs(bpu - (1 << 20) / 11.0) < 1.0


def test_plan_hides_gathers_the_default_policy_exposes():
    """Gathers that take longer than one compute step: one-unit-ahead exposes the difference at every step, the
    planner launches further ahead and hides all but the very first one."""
    from deepspeed_b200.compile import plan_timeline
    n = 8
    ub, su, st, sp = _timeline(n, unit_mb=100.0, step_ms=1.0)
    # 100 MiB at 50 GB/s = 2.1 ms > 1 ms step
    persist, launch, keep, sim = plan_timeline(ub, [0] * n, su, st, sp, 10.0, 50e3, mem_budget=0, max_inflight=0,
                                               persist_headroom=0, reuse_window_us=0.0, safety=1.0)
    assert sim["feasible"] == 1.0 and persist.sum() == 0
    assert sim["default_exposed_us"] > 10e3
    # the communication stream is the bottleneck here (16 gathers x 2.1 ms > 16 ms compute): everything that can be
    # hidden is hidden, i.e. the plan finishes when the serial gather stream does
    total_comm = 2 * n * (10.0 + ub[0] / 50e3)
    assert sim["plan_us"] <= total_comm + st[-1] + 1.0
    assert sim["plan_us"] < sim["default_us"]
    assert launch[n] == -1 and keep[n - 1] == 1  # last forward unit is the first backward unit: never released
    used = [s for s in range(2 * n) if s != n]
    for s in used:
        assert 0 <= launch[s] <= s
    # launch order on the stream follows use order (every rank issues the same collective sequence)
    assert all(launch[a] <= launch[b] for a, b in zip(used[:-1], used[1:]))


def test_plan_respects_memory_budget_and_inflight_cap():
    from deepspeed_b200.compile import plan_timeline
    n = 6
    ub, su, st, sp = _timeline(n, unit_mb=1024.0, step_ms=1.0, peak_gb=10.0)
    budget = int(12.5 * 2**30)  # room for two gathered units next to the 10 GiB peak
    persist, launch, keep, sim = plan_timeline(ub, [0] * n, su, st, sp, 10.0, 50e3, mem_budget=budget,
                                               persist_headroom=0, reuse_window_us=0.0, safety=1.0)
    assert sim["feasible"] == 1.0
    assert sim["plan_peak_bytes"] <= budget
    live = np.zeros(2 * n, dtype=np.int64)
    for s in range(2 * n):
        if launch[s] >= 0:
            live[launch[s]:s + 1 + int(keep[s])] += ub[0]
    assert (live + np.array(sp) <= budget).all() and live.max() == 2 * ub[0]
    # in-flight cap: at most one unit gathered ahead
    persist, launch, keep, sim = plan_timeline(ub, [0] * n, su, st, sp, 10.0, 50e3, mem_budget=0,
                                               max_inflight=ub[0], persist_headroom=0, reuse_window_us=0.0, safety=1.0)
    ahead = np.zeros(2 * n, dtype=np.int64)
    for s in range(2 * n):
        if launch[s] >= 0:
            ahead[launch[s]:s] += ub[0]
    assert ahead.max() <= ub[0]


def test_selective_persistence_prefers_latency_bound_units():
    from deepspeed_b200.compile import plan_timeline
    sizes = [1 << 20, 512 << 20, 2 << 20, 512 << 20, 4 << 20]
    n = len(sizes)
    su = list(range(n)) + list(range(n - 1, -1, -1))
    st = [1000.0] * (2 * n)
    sp = [0] * (2 * n)
    persist, launch, keep, sim = plan_timeline(sizes, [0] * n, su, st, sp, 20.0, 300e3, mem_budget=0,
                                               persist_headroom=8 << 20, reuse_window_us=0.0)
    assert persist.tolist() == [1, 0, 1, 0, 1] and sim["persisted_bytes"] == 7 << 20
    assert all(launch[s] == -1 for s in range(2 * n) if su[s] in (0, 2, 4))
    # derived head-room: budget - peak - the two largest units
    budget = (1024 << 20) + (3 << 20)
    persist, *_ = plan_timeline(sizes, [0] * n, su, st, sp, 20.0, 300e3, mem_budget=budget, persist_headroom=-1)
    assert persist.tolist() == [1, 0, 1, 0, 0]


def test_reuse_window_keeps_last_units_gathered():
    from deepspeed_b200.compile import plan_timeline
    n = 5
    ub, su, st, sp = _timeline(n, unit_mb=100.0, step_ms=1.0)
    persist, launch, keep, sim = plan_timeline(ub, [0] * n, su, st, sp, 10.0, 50e3, persist_headroom=0,
                                               reuse_window_us=2500.0)
    # unit n-1: gap 0; unit n-2: gap 2 ms; unit n-3: gap 4 ms > window
    assert keep.tolist()[:n] == [0, 0, 0, 1, 1]
    assert launch[n] == -1 and launch[n + 1] == -1 and launch[n + 2] >= 0


def _dc_train(enable_dc, persist):
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = SimpleModel(16, 6)
    cfg = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": 3, "unit_max_numel": 300, "stage3_param_persistence_threshold": 0,
                                 "stage3_max_reuse_distance": 0},
           "compile": {"deepcompile": enable_dc, "persist_headroom_bytes": 1 << 30}}
    eng, _, _, _ = ds.initialize(model=model, config=cfg)
    if enable_dc:
        eng.compile(schedule=[(0, ["prefetch", "selective_gather"] if persist else ["prefetch"])])
        assert eng.is_compiled
    co = eng.optimizer.coordinator
    batches = random_batches(16, 4, 16, "cpu")
    losses, gathers = [], []
    for i in range(7):
        g0 = co.gathers_launched
        x, y = batches[2 * i + ds.comm.get_rank()]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
        gathers.append(co.gathers_launched - g0)
    info = {}
    if enable_dc:
        drv = eng._dc_driver
        info = {"planned": drv.plan is not None, "resident": list(drv.plan.resident), "keep": sorted(drv.plan.keep),
                "n": len(co.fwd_order), "lat": drv.profile.comm_latency_us, "bpu": drv.profile.comm_bytes_per_us,
                "step_us": list(drv.profile.step_us)}
    # evaluation forward after training still sees current weights (resident copies were refreshed)
    eng.eval()
    with torch.no_grad():
        ev = float(eng(*batches[15]))
    return losses, gathers, info, ev


@pytest.mark.parametrize("persist", [False, True])
def test_engine_compile_profiles_plans_and_trains_identically(persist):
    base = run_distributed(_dc_train, 2, args=(False, False))
    dc = run_distributed(_dc_train, 2, args=(True, persist))
    for r in range(2):
        assert dc[r][2]["planned"]
        for a, b in zip(base[r][0], dc[r][0]):
            assert abs(a - b) < 1e-6
        assert abs(base[r][3] - dc[r][3]) < 1e-6
    assert dc[0][2]["resident"] == dc[1][2]["resident"] and dc[0][2]["keep"] == dc[1][2]["keep"]
    n = dc[0][2]["n"]
    assert all(t > 0 for t in dc[0][2]["step_us"][:n])
    assert base[0][1][-1] == 2 * n
    if persist:
        # 1 GiB of head-room: every unit is chosen resident and gathered once per optimizer step
        assert len(dc[0][2]["resident"]) == n and dc[0][1][-1] == n
    else:
        assert dc[0][2]["resident"] == [] and n <= dc[0][1][-1] <= 2 * n
