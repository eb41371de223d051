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


def init_z3(engine, backend="inductor", compile_config=None, compile_kwargs=None, schedule=None):
    did = _apply_passes(engine, compile_config, schedule or [(0, "prefetch", {}), (0, "selective_gather", {})])
    log_dist(f"DeepCompile(z3): eager unit scheduler configured: {did}", [0])
    return make_backend(backend, compile_config, compile_kwargs)


def init_z1(engine, backend="inductor", compile_config=None, compile_kwargs=None, schedule=None, use_z2=False):
    did = _apply_passes(engine, compile_config, schedule or [])
    log_dist(f"DeepCompile(z{'2' if use_z2 else '1'}): bucketed reduce on the side stream is the default: {did}", [0])
    return make_backend(backend, compile_config, compile_kwargs)
