"""Multi-process single-host test harness (reference ``tests/unit/common.py:139-520`` DistributedExec /
DistributedTest): forks ``world_size`` local processes, rendezvous on 127.0.0.1, gloo on CPU / NCCL on GPU,
hang detection by timeout, exceptions propagated to the parent."""
import multiprocessing as mp
import os
import socket
import sys
import traceback

import torch

DEFAULT_TIMEOUT = 300
_PRELOADED = False


def free_port():
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world_size, port, fn, args, kwargs, backend, q):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world_size), LOCAL_SIZE=str(world_size),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    try:
        import deepspeed_b200 as ds
        if backend == "nccl":
            torch.cuda.set_device(rank % torch.cuda.device_count())
        ds.init_distributed(dist_backend=backend, verbose=False)
        out = fn(*args, **kwargs)
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
        import pickle
        q.put((rank, None, pickle.dumps(out)))
    except Exception:  # noqa: BLE001
        q.put((rank, traceback.format_exc(), None))


def run_distributed(fn, world_size=2, args=(), kwargs=None, backend=None, timeout=DEFAULT_TIMEOUT):
    """Run ``fn(*args)`` on ``world_size`` ranks; returns list of per-rank return values (must be picklable). The
    rendezvous port is probed and released before the ranks start, so another socket can take it in between: that
    (and only that) failure is retried with a fresh port."""
    for attempt in range(3):
        try:
            return _run_distributed_once(fn, world_size, args, kwargs, backend, timeout)
        except RuntimeError as e:
            if "EADDRINUSE" not in str(e) or attempt == 2:
                raise


def _run_distributed_once(fn, world_size, args, kwargs, backend, timeout):
    kwargs = kwargs or {}
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() and torch.cuda.device_count() >= world_size else "gloo"
    if backend == "gloo":
        # CPU ranks come from a fork server that has torch and the package imported once: a rank starts in ~0.3 s
        # instead of ~6 s of imports, and the server is single threaded, so forking it is safe
        ctx = mp.get_context("forkserver")
        global _PRELOADED
        if not _PRELOADED:
            # torch._dynamo is pulled in lazily by the first torch.optim.Optimizer() and costs ~6 s: preload it too
            ctx.set_forkserver_preload(["torch", "numpy", "torch._dynamo", "torch.distributed.fsdp", "deepspeed_b200",
                                        "deepspeed_b200.runtime.engine", "deepspeed_b200.models"])
            _PRELOADED = True
    else:
        ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=_worker, args=(r, world_size, port, fn, args, kwargs, backend, q), daemon=False)
             for r in range(world_size)]
    for p in procs:
        p.start()
    results, errors = {}, []
    try:
        for _ in range(world_size):
            rank, err, out = q.get(timeout=timeout)
            if err is not None:
                errors.append((rank, err))
                break
            import pickle
            results[rank] = pickle.loads(out)
    except Exception as e:  # queue.Empty -> hang
        import queue as _queue
        if isinstance(e, _queue.Empty):
            errors.append((-1, f"distributed test timed out after {timeout}s (hang?)"))
        else:
            errors.append((-1, "harness error: " + traceback.format_exc()))
    finally:
        for p in procs:
            p.join(timeout=10 if not errors else 1)
            if p.is_alive():
                p.kill()
    if errors:
        raise RuntimeError("\n".join(f"[rank {r}] {e}" for r, e in errors))
    return [results[r] for r in range(world_size)]


class DistributedTest:
    """Subclass with ``world_size = N`` and call ``self.run_ranks(fn, ...)`` from tests."""
    world_size = 2

    def run_ranks(self, fn, *args, world_size=None, **kwargs):
        return run_distributed(fn, world_size or self.world_size, args=args, kwargs=kwargs)
