"""Event counters of the ZeRO-3 parameter coordinator (reference ``runtime/zero/partitioned_param_profiler.py``): how
often and for how many elements parameters were fetched (demanded by a module), prefetched (launched ahead) and
waited for, per pass direction; with timers attached the same events are timed."""
from dataclasses import dataclass

from ...utils.logging import log_dist

FORWARD_FETCH_SUBMIT = "forward_fetch_submit"
FORWARD_FETCH_WAIT = "forward_fetch_wait"
FORWARD_PREFETCH_SUBMIT = "forward_prefetch_submit"
BACKWARD_FETCH_SUBMIT = "backward_fetch_submit"
BACKWARD_FETCH_WAIT = "backward_fetch_wait"
BACKWARD_PREFETCH_SUBMIT = "backward_prefetch_submit"
FORWARD_ALL_GATHER = "forward_all_gather"
BACKWARD_ALL_GATHER = "backward_all_gather"


class PartitionedParameterProfiler:

    @dataclass
    class EventCounter:
        name: str
        count: int
        num_elem: int

        def reset(self):
            self.count = 0
            self.num_elem = 0

        def increment(self, numel):
            self.count += 1
            self.num_elem += numel

    def __init__(self, timers=None):
        self.timers = timers
        self.event_counters = {}

    def reset_events(self):
        for c in self.event_counters.values():
            c.reset()

    def start_event(self, name):
        if name not in self.event_counters:
            self.event_counters[name] = self.EventCounter(name=name, count=0, num_elem=0)
        if self.timers is not None:
            self.timers(name).start()

    def stop_event(self, name, num_elem):
        assert name in self.event_counters, f"unknown event {name}"
        self.event_counters[name].increment(num_elem)
        if self.timers is not None:
            self.timers(name).stop()

    def count(self, name, num_elem):
        """start + stop for an instantaneous event."""
        self.start_event(name)
        self.stop_event(name, num_elem)

    def log_events(self):
        for c in self.event_counters.values():
            log_dist(f"{c.name}: count = {c.count}, numel = {c.num_elem}", ranks=[0])
        if self.timers is not None and self.event_counters:
            self.timers.log(names=list(self.event_counters.keys()))
