"""Entry points with the reference's names (``compile/backend.py``, ``init_z1_and_2.py``, ``init_z3.py``)."""
import torch

from ..utils.logging import log_dist


def is_deepcompile_supported():
    """The scheduling DeepCompile adds is native to the eager engine here; nothing extra needs compiling."""
    return True


def _eager_backend(gm, example_inputs):
    return gm.forward


def make_backend(backend="inductor", compile_config=None, compile_kwargs=None, free_activation=False, debug_log=False):
    """A ``torch.compile`` backend. ``"eager"``/``None`` returns the graph unchanged; ``"inductor"`` defers to torch's
    own backend for the elementwise glue the user may still want compiled (never required by the framework)."""
    if backend in (None, "eager"):
        return _eager_backend
    if backend == "inductor":
        from torch._dynamo.backends.inductor import inductor
        return inductor
    raise ValueError(f"Unsupported backend {backend}")


def _apply_passes(engine, compile_config, schedule):
    z = engine.optimizer
    did = []
    coord = getattr(z, "coordinator", None)
    for step, name, kwargs in (schedule or []):
        if name == "prefetch" and coord is not None:
            coord.prefetch = True
            did.append("prefetch (one unit ahead on the all-gather stream)")
        elif name == "selective_gather" and coord is not None:
            budget = kwargs.get("persistent_bytes") if kwargs else None
            n = coord.mark_persistent_by_budget(budget) if hasattr(coord, "mark_persistent_by_budget") else 0
            did.append(f"selective_gather ({n} units kept resident)")
        elif name in ("offload_adam_states", "offload_adam_states_sync"):
            z.offload_states(include=["optim_states"])
            did.append("offload_adam_states")
        elif name == "offload_parameters":
            did.append("offload_parameters (use zero_optimization.offload_param)")
    return did


_CUSTOM_PASSES = {}  # name -> fn(profile, plan) -> plan | None


def register_compile_pass(pass_name, pass_fn, contract=None):
    """Reference ``compile/passes/__init__.py:register_compile_pass``. Here a pass edits the :class:`UnitPlan` that the
    built-in planner produced (launch points, kept units, resident units) given the :class:`UnitProfile`."""
    if not callable(pass_fn):
        raise TypeError("pass_fn must be callable: pass_fn(profile, plan) -> plan or None")
    _CUSTOM_PASSES[pass_name] = pass_fn


class UnitScheduleDriver:
    """Profile -> plan -> execute loop of the ZeRO-3 unit executor (the reference's ``launch_compile_passes`` schedule,
    ``compile/backend.py:200-280``, at unit granularity): the first optimizer step traces the unit order, the next
    ``profile_steps`` run under :class:`UnitProfiler`, then the native planner (``csrc/cpu/dc_planner.cpp``) produces the
    gather schedule the coordinator follows from there on. ``passes`` selects what the plan may do:
    ``prefetch`` (launch points) and ``selective_gather`` (resident units)."""

    def __init__(self, engine, passes=("prefetch", "selective_gather"), profile_steps=2, mem_margin=0.1,
                 max_inflight=int(4e9), persist_headroom=None, debug_log=False):
        from .profiler import UnitProfiler
        self.engine = engine
        self.coord = engine.optimizer.coordinator
        self.passes = set(passes)
        self.profile_steps = max(1, int(profile_steps))
        self.mem_margin, self.max_inflight, self.debug_log = mem_margin, max_inflight, debug_log
        self.persist_headroom = persist_headroom  # bytes for resident units; None: derived from the profiled peak
        self.profiler = UnitProfiler(self.coord)
        self.plan = None
        self.profile = None
        self.pass_times = {}
        self._seen = 0

    def tick(self):
        """Called after every optimizer step."""
        if self.plan is not None:
            return
        c = self.coord
        if c._prof is None:
            if c._trace_done:
                c._prof = self.profiler  # profile from the next step on (the trace step itself is not representative)
            return
        self._seen += 1
        if self._seen < self.profile_steps:
            return
        import time
        from .planner import plan_units
        c._prof = None
        t0 = time.perf_counter()
        self.profile = self.profiler.result()
        t1 = time.perf_counter()
        self.plan = plan_units(self.profile, mem_margin=self.mem_margin, max_inflight=self.max_inflight,
                               persist="selective_gather" in self.passes, persist_headroom=self.persist_headroom)
        if "prefetch" not in self.passes:
            self.plan.launch_at = {}
            self.plan.keep = set()
        self.pass_times = {"profile": t1 - t0, "plan": time.perf_counter() - t1}
        for name in self.passes:
            fn = _CUSTOM_PASSES.get(name)
            if fn is not None:
                t2 = time.perf_counter()
                out = fn(self.profile, self.plan)
                if out is not None:
                    self.plan = out
                self.pass_times[name] = time.perf_counter() - t2
        c.set_plan(self.plan)
        pr = self.plan.predicted
        log_dist(f"DeepCompile(z3): unit plan installed: {len(self.plan.resident)} resident units "
                 f"({pr['persisted_bytes'] / 2**20:.0f} MiB), {len(self.plan.keep)} kept forward->backward, predicted "
                 f"exposed gather wait {pr['plan_exposed_us'] / 1e3:.2f} ms (one-ahead policy "
                 f"{pr['default_exposed_us'] / 1e3:.2f} ms), comm model {self.profile.comm_latency_us:.1f} us + "
                 f"{self.profile.comm_bytes_per_us / 1e3:.1f} GB/s", [0])


def init_z3(engine, backend="inductor", compile_config=None, compile_kwargs=None, schedule=None):
    """Reference ``compile/init_z3.py:init_z3``. ``schedule``: ``[(step, [pass, ...]), ...]`` (names, or callables whose
    ``__name__`` / ``NAME`` is the pass name); the union of the passes configures the unit plan."""
    names = []
    for entry in (schedule or []):
        ps = entry[1] if isinstance(entry[1], (list, tuple)) else [entry[1]]
        for p in ps:
            names.append(p if isinstance(p, str) else getattr(p, "NAME", getattr(p, "__name__", str(p))))
    if not names:
        names = ["prefetch", "selective_gather"]
    did = _apply_passes(engine, compile_config, [(0, n, {}) for n in names if n.startswith("offload")])
    coord = getattr(engine.optimizer, "coordinator", None)
    if coord is not None:
        unit_passes = [n for n in names if n in ("prefetch", "selective_gather") or n in _CUSTOM_PASSES]
        extra = lambda k, d: getattr(compile_config, k, d) if compile_config is not None else d  # noqa: E731
        engine._dc_driver = UnitScheduleDriver(engine, passes=unit_passes, profile_steps=extra("profile_steps", 2),
                                               mem_margin=extra("mem_margin", 0.1),
                                               max_inflight=int(extra("max_buffered_size", 4e9)),
                                               persist_headroom=extra("persist_headroom_bytes", None),
                                               debug_log=bool(extra("debug_log", False)))
        did.append(f"unit plan after profiling ({', '.join(unit_passes) or 'none'})")
    log_dist(f"DeepCompile(z3): {did}", [0])
    return make_backend(backend, compile_config, compile_kwargs)


def init_z1(engine, backend="inductor", compile_config=None, compile_kwargs=None, schedule=None, use_z2=False):
    did = _apply_passes(engine, compile_config, schedule or [])
    log_dist(f"DeepCompile(z{'2' if use_z2 else '1'}): bucketed reduce on the side stream is the default: {did}", [0])
    return make_backend(backend, compile_config, compile_kwargs)
