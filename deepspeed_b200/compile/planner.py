"""Unit-timeline planner: Python face of ``csrc/cpu/dc_planner.cpp``.

Reference: the DeepCompile passes ``compile/passes/prefetch.py`` (move all-gathers early inside the profiled memory
budget), ``compile/passes/selective_gather.py`` (keep the parameters with the best gather-time-per-byte resident) and
the native executor ``csrc/compile/z3.cpp``. Here the executor is the ZeRO-3 unit coordinator
(``runtime/zero/coordinator.py``); the plan tells it at which timeline step to launch each gather, which units stay
gathered from forward into backward, and which stay resident across micro-steps.

Timeline convention: ``n`` units in first-forward order give ``2n`` steps - step ``p`` is the forward of position ``p``,
step ``2n-1-p`` its backward.
"""
from dataclasses import dataclass, field
from typing import Dict, List, Set

import numpy as np

from ..ops import cpu_lib


@dataclass
class UnitProfile:
    """One profiled training step at unit granularity (all ranks hold the same, all-reduced, numbers)."""
    unit_ids: List[int]  # unit id per forward position
    unit_bytes: List[int]  # gathered bytes per forward position
    step_us: List[float]  # 2n compute times (forward steps then backward steps)
    step_peak: List[int]  # 2n peak device bytes inside the step, gathered unit buffers excluded
    comm_latency_us: float = 20.0
    comm_bytes_per_us: float = 300e3  # 300 GB/s
    total_memory: int = 0  # device capacity in bytes (0: unlimited)


@dataclass
class UnitPlan:
    launch_at: Dict[int, List[int]] = field(default_factory=dict)  # step -> steps whose gather is issued there
    keep: Set[int] = field(default_factory=set)  # forward positions not released after use
    resident: List[int] = field(default_factory=list)  # unit ids that stay gathered across micro-steps
    predicted: Dict[str, float] = field(default_factory=dict)


def fit_comm(nbytes, micros):
    """Least-squares ``time = latency + bytes / bandwidth``; returns (latency_us, bytes_per_us)."""
    n = len(nbytes)
    b = np.ascontiguousarray(nbytes, dtype=np.int64)
    t = np.ascontiguousarray(micros, dtype=np.float64)
    out = np.zeros(2, dtype=np.float64)
    rc = cpu_lib.lib().dsb200_dc_fit_comm(n, b.ctypes.data, t.ctypes.data, out.ctypes.data)
    if rc != 0:
        raise ValueError("fit_comm needs at least one sample")
    return float(out[0]), float(out[1])


def plan_timeline(unit_bytes, persistent, step_unit, step_us, step_peak, latency_us, bytes_per_us, mem_budget=0,
                  max_inflight=0, persist_headroom=-1, reuse_window_us=-1.0, safety=1.5):
    """Raw planner call on an arbitrary timeline. Returns (persist[n_units], launch_step[n_steps], keep_after[n_steps],
    sim dict)."""
    nu, ns = len(unit_bytes), len(step_unit)
    ub = np.ascontiguousarray(unit_bytes, dtype=np.int64)
    pin = np.ascontiguousarray(persistent, dtype=np.uint8)
    su = np.ascontiguousarray(step_unit, dtype=np.int32)
    st = np.ascontiguousarray(step_us, dtype=np.float64)
    sp = np.ascontiguousarray(step_peak, dtype=np.int64)
    pout = np.zeros(nu, dtype=np.uint8)
    launch = np.zeros(ns, dtype=np.int32)
    keep = np.zeros(ns, dtype=np.int32)
    sim = np.zeros(8, dtype=np.float64)
    rc = cpu_lib.lib().dsb200_dc_plan(nu, ub.ctypes.data, pin.ctypes.data, ns, su.ctypes.data, st.ctypes.data,
                                      sp.ctypes.data, float(latency_us), float(bytes_per_us), int(mem_budget),
                                      int(max_inflight), int(persist_headroom), float(reuse_window_us), float(safety),
                                      pout.ctypes.data, launch.ctypes.data, keep.ctypes.data, sim.ctypes.data)
    if rc != 0:
        raise ValueError(f"dsb200_dc_plan failed ({rc})")
    keys = ["plan_us", "plan_exposed_us", "plan_peak_bytes", "default_us", "default_exposed_us", "default_peak_bytes",
            "persisted_bytes", "feasible"]
    return pout, launch, keep, dict(zip(keys, sim.tolist()))


def plan_units(profile: UnitProfile, mem_margin=0.1, max_inflight=int(4e9), persist=True, persist_headroom=None,
               reuse_window_us=-1.0, safety=1.5) -> UnitPlan:
    """Plan for a profiled step. ``mem_margin``: share of the device memory left unused (reference MARGIN 0.1);
    ``max_inflight``: bytes gathered ahead of use at any time (reference MAX_BUFFERED_SIZE 4e9); ``persist_headroom``:
    bytes that may go to resident units (None: budget - profiled peak - two largest units; nothing when the device
    memory is unknown)."""
    n = len(profile.unit_ids)
    assert len(profile.step_us) == 2 * n and len(profile.step_peak) == 2 * n
    step_unit = list(range(n)) + list(range(n - 1, -1, -1))
    budget = int(profile.total_memory * (1 - mem_margin)) if profile.total_memory else 0
    pout, launch, keep, sim = plan_timeline(profile.unit_bytes, [0] * n, step_unit, profile.step_us,
                                            profile.step_peak, profile.comm_latency_us, profile.comm_bytes_per_us,
                                            mem_budget=budget, max_inflight=max_inflight,
                                            persist_headroom=(0 if not persist else -1 if persist_headroom is None
                                                              else int(persist_headroom)),
                                            reuse_window_us=reuse_window_us,
                                            safety=safety)
    plan = UnitPlan(predicted=sim)
    for s in range(2 * n):
        if launch[s] >= 0:
            plan.launch_at.setdefault(int(launch[s]), []).append(s)
        if keep[s] and s < n:
            plan.keep.add(s)
    plan.resident = [profile.unit_ids[p] for p in range(n) if pout[p]]
    return plan
