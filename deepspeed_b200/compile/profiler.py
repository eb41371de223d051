"""Unit-granular step profiler feeding the planner.

Reference: ``compile/profilers/graph_profile.py`` (per-node device time, wall time, allocated / peak memory of the FX
graph) and ``compile/profilers/comm_profile.py`` (all-gather time model). Here the "nodes" are the ZeRO-3 flat units:
the coordinator reports when a unit's forward / backward starts, when a pass ends and when a gather is launched; on a
GPU the marks are CUDA events on the launching stream and the memory numbers come from the caching allocator, on CPU
they are wall-clock stamps.
"""
import time

import torch

from .. import comm as dist
from .planner import UnitProfile, fit_comm


class UnitProfiler:

    def __init__(self, coordinator):
        self.c = coordinator
        self.cuda = torch.cuda.is_available() and any(u.device.type == "cuda" for u in coordinator.z.units)
        self.passes = []  # finished passes: list of marks
        self._marks = []
        self.gathers = []  # (bytes, begin, end)

    # ---- called by the coordinator ------------------------------------------------------------
    def _stamp(self, stream=None):
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(stream if stream is not None else torch.cuda.current_stream())
            return ev
        return time.perf_counter()

    def _gathered_bytes(self):
        return sum(u.numel_padded * torch.empty(0, dtype=u.dtype).element_size() for u in self.c.z.units
                   if not u.persistent and (u.gathered or u.gather_work) and u.world > 1)

    def step_begin(self, phase, units):
        if self._marks and self.cuda:
            m = self._marks[-1]
            m["peak"] = torch.cuda.max_memory_allocated() - m["gathered_peak"]
        mark = {"phase": phase, "uids": [u.uid for u in units], "t": self._stamp(), "peak": 0, "gathered_peak": 0}
        if self.cuda:
            torch.cuda.reset_peak_memory_stats()
            mark["gathered_peak"] = self._gathered_bytes()
        self._marks.append(mark)

    def pass_end(self):
        if not self._marks:
            return
        if self.cuda:
            m = self._marks[-1]
            m["peak"] = torch.cuda.max_memory_allocated() - m["gathered_peak"]
        self._marks.append({"phase": "end", "uids": [], "t": self._stamp(), "peak": 0})
        self.passes.append(self._marks)
        self._marks = []

    def gather_begin(self, u, stream):
        return (u.numel_padded * torch.empty(0, dtype=u.dtype).element_size(), self._stamp(stream))

    def gather_end(self, tok, stream):
        self.gathers.append((tok[0], tok[1], self._stamp(stream)))

    # ---- result -------------------------------------------------------------------------------
    def _delta_us(self, a, b):
        return a.elapsed_time(b) * 1e3 if self.cuda else (b - a) * 1e6

    def result(self, group=None) -> UnitProfile:
        """Profile of the last complete pass, all-reduced (max) over the ranks so every rank plans identically."""
        if self.cuda:
            torch.cuda.synchronize()
        order = self.c.fwd_order
        n = len(order)
        pos_of = {uid: p for p, uid in enumerate(order)}
        marks = None
        for cand in reversed(self.passes):  # the last pass that saw every unit forward and backward
            seen_f = {uid for m in cand if m["phase"] == "fwd" for uid in m["uids"]}
            seen_b = {uid for m in cand if m["phase"] == "bwd" for uid in m["uids"]}
            if set(order) <= seen_f and set(order) <= seen_b:
                marks = cand
                break
        if marks is None:
            raise RuntimeError("UnitProfiler: no complete forward+backward pass recorded")
        step_us = [0.0] * (2 * n)
        step_peak = [0] * (2 * n)
        for m, nxt in zip(marks[:-1], marks[1:]):
            steps = [pos_of[uid] if m["phase"] == "fwd" else 2 * n - 1 - pos_of[uid] for uid in m["uids"]
                     if uid in pos_of]
            if not steps:
                continue
            s = min(steps)  # several units used by one module: the time goes to the first step in timeline order
            step_us[s] += self._delta_us(m["t"], nxt["t"])
            for k in steps:
                step_peak[k] = max(step_peak[k], int(m["peak"]))
        esz = {u.uid: torch.empty(0, dtype=u.dtype).element_size() for u in self.c.z.units}
        unit_bytes = [self.c.z.units[uid].numel_padded * esz[uid] for uid in order]
        lat, bpu = 20.0, 300e3
        samples = [(b, self._delta_us(t0, t1)) for b, t0, t1 in self.gathers]
        samples = [(b, t) for b, t in samples if t > 0]
        if samples:
            lat, bpu = fit_comm([b for b, _ in samples], [t for _, t in samples])
        total = torch.cuda.get_device_properties(0).total_memory if self.cuda else 0
        vals = torch.tensor(step_us + [float(p) for p in step_peak] + [lat, -bpu, -float(total)], dtype=torch.float64)
        if dist.is_initialized() and dist.get_world_size() > 1:
            dev = torch.device("cuda", torch.cuda.current_device()) if self.cuda else torch.device("cpu")
            vals = vals.to(dev)
            dist.all_reduce(vals, op=dist.ReduceOp.MAX, group=group)  # slowest time, largest peak, lowest bandwidth
            vals = vals.cpu()
        vals = vals.tolist()
        return UnitProfile(unit_ids=list(order), unit_bytes=unit_bytes, step_us=vals[:2 * n],
                           step_peak=[int(v) for v in vals[2 * n:4 * n]], comm_latency_us=vals[4 * n],
                           comm_bytes_per_us=-vals[4 * n + 1], total_memory=int(-vals[4 * n + 2]))
