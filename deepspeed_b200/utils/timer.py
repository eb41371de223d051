"""Timers: CUDA-event wall-clock timers, throughput timer.

Reference behaviour: ``deepspeed/utils/timer.py:32-162`` (SynchronizedWallClockTimer with named timers,
event based, ``log(names, normalizer, reset, memory_breakdown, ranks)``), ``:199-294`` (ThroughputTimer
skipping ``start_step`` warm-up steps, samples/sec every ``steps_per_output``).
"""
import time

import torch

from .logging import log_dist, logger

FORWARD_MICRO_TIMER = "fwd_microstep"
FORWARD_GLOBAL_TIMER = "fwd"
BACKWARD_MICRO_TIMER = "bwd_microstep"
BACKWARD_GLOBAL_TIMER = "bwd"
BACKWARD_INNER_MICRO_TIMER = "bwd_inner_microstep"
BACKWARD_INNER_GLOBAL_TIMER = "bwd_inner"
BACKWARD_REDUCE_MICRO_TIMER = "bwd_allreduce_microstep"
BACKWARD_REDUCE_GLOBAL_TIMER = "bwd_allreduce"
STEP_MICRO_TIMER = "step_microstep"
STEP_GLOBAL_TIMER = "step"
TIME_EPSILON = 1e-6


def _cuda():
    return torch.cuda.is_available()


class _Timer:
    """One named timer. On CUDA records event pairs and resolves them lazily (no sync on start/stop)."""

    def __init__(self, name):
        self.name_ = name
        self.started_ = False
        self.pairs = []
        self.start_event = None
        self.elapsed_records = 0.0
        self.count = 0
        self._t0 = 0.0

    def start(self):
        assert not self.started_, f"{self.name_} timer has already been started"
        if _cuda():
            self.start_event = torch.cuda.Event(enable_timing=True)
            self.start_event.record()
        else:
            self._t0 = time.perf_counter()
        self.started_ = True

    def stop(self, reset=False, record=False):
        assert self.started_, f"{self.name_} timer is not started"
        if _cuda():
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            self.pairs.append((self.start_event, end))
            self.start_event = None
        else:
            self.elapsed_records += (time.perf_counter() - self._t0) * 1e3
        self.count += 1
        self.started_ = False

    def _resolve(self):
        if self.pairs:
            self.pairs[-1][1].synchronize()
            self.elapsed_records += sum(s.elapsed_time(e) for s, e in self.pairs)
            self.pairs = []

    def elapsed(self, reset=True):
        """Milliseconds accumulated since last reset."""
        was = self.started_
        if was:
            self.stop()
        self._resolve()
        val = self.elapsed_records
        if reset:
            self.reset()
        if was:
            self.start()
        return val

    def mean(self):
        self._resolve()
        return self.elapsed_records / max(1, self.count)

    def reset(self):
        self.started_ = False
        self.pairs = []
        self.elapsed_records = 0.0
        self.count = 0


class SynchronizedWallClockTimer:
    Timer = _Timer

    def __init__(self):
        self.timers = {}

    def get_timers(self):
        return self.timers

    def __call__(self, name):
        if name not in self.timers:
            self.timers[name] = _Timer(name)
        return self.timers[name]

    def has(self, name):
        return name in self.timers

    @staticmethod
    def memory_usage():
        if not _cuda():
            return ""
        gb = 1024**3
        return (f" | mem_allocated: {torch.cuda.memory_allocated() / gb:.4f} GB"
                f" max_mem_allocated: {torch.cuda.max_memory_allocated() / gb:.4f} GB"
                f" cache_allocated: {torch.cuda.memory_reserved() / gb:.4f} GB"
                f" max_cache_allocated: {torch.cuda.max_memory_reserved() / gb:.4f} GB")

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None):
        assert normalizer > 0.0
        s = "time (ms)"
        for n in names:
            if n in self.timers:
                s += f" | {n}: {self.timers[n].elapsed(reset=reset) / normalizer:.2f}"
        if memory_breakdown:
            s += self.memory_usage()
        log_dist(s, ranks=ranks or [0])

    def get_mean(self, names, normalizer=1.0, reset=True):
        means = {}
        for n in names:
            if n in self.timers:
                means[n] = self.timers[n].mean() * 1000.0 / normalizer
                if reset:
                    self.timers[n].reset()
        return means


class NoopTimer:

    class _T:

        def start(self):
            ...

        def reset(self):
            ...

        def stop(self, **kwargs):
            ...

        def elapsed(self, **kwargs):
            return 0

        def mean(self):
            return 0

    def __init__(self):
        self.timer = self._T()

    def __call__(self, name):
        return self.timer

    def get_timers(self):
        return {}

    def log(self, *a, **k):
        ...

    def get_mean(self, *a, **k):
        return {}

    def has(self, name):
        return False


class ThroughputTimer:

    def __init__(self, config=None, batch_size=1, start_step=2, steps_per_output=None, monitor_memory=False,
                 logging_fn=None):
        self.config = config
        self.enabled = True if config is None else getattr(config, "enabled", True)
        self.synchronized = True if config is None else getattr(config, "synchronized", True)
        self.start_time = 0
        self.end_time = 0
        self.started = False
        self.batch_size = 1 if batch_size is None else batch_size
        self.start_step = start_step
        self.epoch_count = 0
        self.micro_step_count = 0
        self.global_step_count = 0
        self.total_elapsed_time = 0
        self.step_elapsed_time = 0
        self.steps_per_output = steps_per_output
        self.monitor_memory = monitor_memory
        self.logging = logging_fn or logger.info
        self.initialized = False

    def update_epoch_count(self):
        self.epoch_count += 1
        self.micro_step_count = 0

    def _init_timer(self):
        self.initialized = True

    def _sync(self):
        if self.synchronized and _cuda():
            torch.cuda.synchronize()

    def start(self):
        if not self.enabled:
            return
        self._init_timer()
        self.started = True
        if self.global_step_count >= self.start_step:
            self._sync()
            self.start_time = time.time()

    def stop(self, global_step=False, report_speed=True):
        if not self.enabled or not self.started:
            return
        self.started = False
        self.micro_step_count += 1
        if global_step:
            self.global_step_count += 1
        if self.start_time > 0:
            self._sync()
            self.end_time = time.time()
            duration = self.end_time - self.start_time
            self.total_elapsed_time += duration
            self.step_elapsed_time += duration
            if global_step:
                if report_speed and self.steps_per_output and self.global_step_count % self.steps_per_output == 0:
                    self.logging(
                        "epoch={}/micro_step={}/global_step={}, RunningAvgSamplesPerSec={}, CurrSamplesPerSec={}, "
                        "MemAllocated={}GB, MaxMemAllocated={}GB".format(
                            self.epoch_count, self.micro_step_count, self.global_step_count, self.avg_samples_per_sec(),
                            self.batch_size / (self.step_elapsed_time + TIME_EPSILON),
                            round(torch.cuda.memory_allocated() / 1024**3, 2) if _cuda() else 0,
                            round(torch.cuda.max_memory_allocated() / 1024**3, 2) if _cuda() else 0))
                self.step_elapsed_time = 0

    def avg_samples_per_sec(self):
        if self.global_step_count > 0:
            total_step_offset = self.global_step_count - self.start_step
            avg_time_per_step = self.total_elapsed_time / max(1, total_step_offset)
            return self.batch_size / (avg_time_per_step + TIME_EPSILON)
        return float("-inf")


def trim_mean(data, trim_percent):
    """Mean after trimming ``trim_percent`` from both ends."""
    assert 0.0 <= trim_percent <= 1.0
    n = len(data)
    if n == 0:
        return 0
    data = sorted(data)
    k = int(round(n * trim_percent))
    return sum(data[k:n - k]) / max(1, n - 2 * k)


class CudaEventTimer:
    """Elapsed device time between two recorded events (reference ``utils/timer.py:32``)."""

    def __init__(self, start_event, end_event):
        self.start_event, self.end_event = start_event, end_event

    def get_elapsed_msec(self):
        import torch
        torch.cuda.current_stream().wait_event(self.end_event)
        self.end_event.synchronize()
        return self.start_event.elapsed_time(self.end_event)
