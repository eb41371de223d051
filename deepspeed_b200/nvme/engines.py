"""I/O engines of the DeepNVMe benchmark (reference ``deepspeed/nvme``: ``ds_aio_basic.py`` ``AIOBasic_Engine``,
``ds_aio_handle.py`` ``AIOHandle_Engine``, ``torch_io.py`` ``TorchIO_Engine``, ``torch_fastio_engine.py``
``Torch_FastIO_Engine``, ``ds_aio_job.py`` ``Job`` / ``run_job``, ``perf_run_sweep.py`` ``SweepConfig``).

Every engine moves ``io_size`` bytes between a host (pinned) or device buffer and one file per thread id and reports
what it did through ``ctxt``: ``elapsed_sec`` per call and ``num_bytes``. ``run(engine_cls, args)`` drives the loops
and turns that into GB/s. ``args`` is any namespace with ``folder``, ``io_size``, ``block_size``, ``queue_depth``,
``intra_op_parallelism``, ``single_submit``, ``overlap_events``, ``loops``, ``use_gpu`` (others optional)."""
import os
import shlex
import subprocess
import time

import torch

from ..ops.aio import aio_handle


def _get(args, name, default):
    return getattr(args, name, default)


def _file(args, tid, read_op):
    folder = args.folder if isinstance(args.folder, str) else args.folder[tid % len(args.folder)]
    os.makedirs(folder, exist_ok=True)
    return os.path.join(folder, f"_aio_{'read' if read_op else 'write'}_{int(args.io_size)}_bytes_{tid}.pt")


def _buffer(args):
    n = int(args.io_size)
    if _get(args, "use_gpu", False) and torch.cuda.is_available():
        return torch.empty(n, dtype=torch.uint8, device="cuda")
    return torch.empty(n, dtype=torch.uint8, pin_memory=torch.cuda.is_available())


def _handle(args):
    return aio_handle(int(_get(args, "block_size", 1 << 20)), int(_get(args, "queue_depth", 8)),
                      bool(_get(args, "single_submit", False)), bool(_get(args, "overlap_events", True)),
                      int(_get(args, "intra_op_parallelism", 1)))


class _Engine:

    def __init__(self, args, tid, read_op):
        self.args, self.tid, self.read_op = args, tid, read_op
        self.ctxt = {"file": _file(args, tid, read_op), "num_bytes": int(args.io_size), "elapsed_sec": []}
        self.ctxt["buffer"] = _buffer(args)
        if read_op and (not os.path.exists(self.ctxt["file"]) or
                        os.path.getsize(self.ctxt["file"]) < self.ctxt["num_bytes"]):
            with open(self.ctxt["file"], "wb") as f:
                f.write(os.urandom(self.ctxt["num_bytes"]))

    def _timed(self, fn):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.ctxt["elapsed_sec"].append(time.perf_counter() - t0)

    def fini(self):
        self.ctxt["buffer"] = None
        if not self.read_op and not _get(self.args, "keep_files", False) and os.path.exists(self.ctxt["file"]):
            os.remove(self.ctxt["file"])

    def read(self, args=None, tid=None, loop_id=0):
        raise NotImplementedError

    def write(self, args=None, tid=None, loop_id=0):
        raise NotImplementedError


class AIOBasic_Engine(_Engine):
    """One blocking request per call: a fresh handle with a single I/O thread, submit and wait."""

    def _h(self):
        a = self.args
        return aio_handle(int(_get(a, "block_size", 1 << 20)), int(_get(a, "queue_depth", 8)),
                          bool(_get(a, "single_submit", False)), bool(_get(a, "overlap_events", True)), 1)

    def _host(self):
        b = self.ctxt["buffer"]
        return b if not b.is_cuda else torch.empty(b.numel(), dtype=torch.uint8, pin_memory=True)

    def read(self, args=None, tid=None, loop_id=0):
        host = self._host()

        def go():
            self._h().sync_pread(host, self.ctxt["file"])
            if host is not self.ctxt["buffer"]:
                self.ctxt["buffer"].copy_(host)

        self._timed(go)

    def write(self, args=None, tid=None, loop_id=0):
        host = self._host()

        def go():
            if host is not self.ctxt["buffer"]:
                host.copy_(self.ctxt["buffer"])
            self._h().sync_pwrite(host, self.ctxt["file"])

        self._timed(go)


class AIOHandle_Engine(_Engine):
    """A persistent handle with ``intra_op_parallelism`` I/O threads; requests are asynchronous and a bounce buffer
    stands between the file and a device buffer."""

    def __init__(self, args, tid, read_op):
        super().__init__(args, tid, read_op)
        self.ctxt["handle"] = _handle(args)
        b = self.ctxt["buffer"]
        self.ctxt["bounce"] = torch.empty(b.numel(), dtype=torch.uint8, pin_memory=True) if b.is_cuda else None

    def read(self, args=None, tid=None, loop_id=0):
        h, b, bounce = self.ctxt["handle"], self.ctxt["buffer"], self.ctxt["bounce"]

        def go():
            h.async_pread(bounce if bounce is not None else b, self.ctxt["file"])
            h.wait()
            if bounce is not None:
                b.copy_(bounce, non_blocking=True)

        self._timed(go)

    def write(self, args=None, tid=None, loop_id=0):
        h, b, bounce = self.ctxt["handle"], self.ctxt["buffer"], self.ctxt["bounce"]

        def go():
            if bounce is not None:
                bounce.copy_(b)
            h.async_pwrite(bounce if bounce is not None else b, self.ctxt["file"])
            h.wait()

        self._timed(go)

    def fini(self):
        self.ctxt["handle"] = self.ctxt["bounce"] = None
        super().fini()


class TorchIO_Engine(_Engine):
    """``torch.save`` / ``torch.load`` on the same bytes: the baseline the native path is compared with."""

    def __init__(self, args, tid, read_op):
        super().__init__(args, tid, read_op)
        if read_op:
            torch.save(self.ctxt["buffer"].cpu(), self.ctxt["file"])

    def read(self, args=None, tid=None, loop_id=0):
        self._timed(lambda: self.ctxt["buffer"].copy_(torch.load(self.ctxt["file"], weights_only=False)))

    def write(self, args=None, tid=None, loop_id=0):
        self._timed(lambda: torch.save(self.ctxt["buffer"], self.ctxt["file"]))


class Torch_FastIO_Engine(_Engine):
    """``torch.save`` through :class:`deepspeed_b200.io.FastFileWriter` (pinned double buffer + native async writes):
    the checkpoint write path."""

    def __init__(self, args, tid, read_op):
        assert not read_op, "the fast writer engine measures writes"
        super().__init__(args, tid, read_op)
        from ..io import FastFileWriterConfig
        self.ctxt["config"] = FastFileWriterConfig(io_buffer_size=int(_get(args, "io_buffer_size", 64 << 20)),
                                                   double_buffer=bool(_get(args, "io_buffer_double", True)),
                                                   num_threads=int(_get(args, "intra_op_parallelism", 4)),
                                                   block_size=int(_get(args, "block_size", 1 << 20)),
                                                   queue_depth=int(_get(args, "queue_depth", 8)))

    def write(self, args=None, tid=None, loop_id=0):
        from ..io import FastFileWriter

        def go():
            w = FastFileWriter(self.ctxt["file"], self.ctxt["config"])
            torch.save(self.ctxt["buffer"], w, _use_new_zipfile_serialization=False)
            w.close()

        self._timed(go)


def run(engine_cls, args, tid=0, read_op=True):
    """``loops`` timed operations (the first ``warmup_loops`` are dropped); returns the summary dict."""
    eng = engine_cls(args, tid, read_op)
    loops, warm = int(_get(args, "loops", 3)), int(_get(args, "warmup_loops", 1))
    for i in range(loops + warm):
        (eng.read if read_op else eng.write)(args, tid, i)
    secs = eng.ctxt["elapsed_sec"][warm:]
    n = eng.ctxt["num_bytes"]
    eng.fini()
    best = min(secs)
    return {"engine": engine_cls.__name__, "op": "read" if read_op else "write", "num_bytes": n, "loops": loops,
            "sec_best": best, "sec_mean": sum(secs) / len(secs), "gbps_best": n / best / 1e9}


class Job:
    """A command line with the file its output goes to (reference ``ds_aio_job.py:12``)."""

    def __init__(self, cmd_line, output_file=None, work_dir=None):
        self.cmd_line, self.output_file, self.work_dir = cmd_line, output_file, work_dir
        self.output_fd = None

    def cmd(self):
        return self.cmd_line

    def get_stdout(self):
        return self.output_fd

    def get_stderr(self):
        return self.output_fd

    def get_cwd(self):
        return self.work_dir

    def open_output_file(self):
        if self.output_file is not None:
            self.output_fd = open(self.output_file, "w")

    def close_output_file(self):
        if self.output_fd is not None:
            self.output_fd.close()
            self.output_fd = None


def run_job(job, verbose=False):
    argv = shlex.split(" ".join(job.cmd()) if not isinstance(job.cmd(), str) else job.cmd())
    if verbose:
        print(f"args = {argv}")
    job.open_output_file()
    try:
        proc = subprocess.run(argv, stdout=job.get_stdout(), stderr=job.get_stderr(), cwd=job.get_cwd())
    finally:
        job.close_output_file()
    assert proc.returncode == 0, f"This command failed: {job.cmd()}"


class SweepConfig:
    """What ``ds_nvme_tune`` was asked to do (reference ``perf_run_sweep.py:52``)."""

    def __init__(self, args):
        self.folder_to_device_mapping = list(_get(args, "nvme_dir", []))
        self.search_space = None
        from . import DEFAULT_SWEEP_CONFIG
        cfg = _get(args, "sweep_config", None)
        if cfg is None:
            self.search_space = dict(DEFAULT_SWEEP_CONFIG)
        elif isinstance(cfg, dict):
            self.search_space = cfg
        else:
            import json
            with open(cfg) as f:
                self.search_space = json.load(f)
        self.search_space.setdefault("gpu", [bool(_get(args, "gpu", False))])
        self.search_space.setdefault("gds", [bool(_get(args, "gds", False))])
        self.read = not bool(_get(args, "no_read", False))
        self.write = not bool(_get(args, "no_write", False))
        self.flush_cache = bool(_get(args, "flush_page_cache", False))
        self.log_dir = _get(args, "log_dir", "_aio_bench_logs")
        self.verbose = bool(_get(args, "verbose", False))
        self.other_options = f"--loops {_get(args, 'loops', 1)} --io_size {_get(args, 'io_size', '400M')}"
        if _get(args, "sequential_requests", False):
            self.other_options += " --sequential_requests"

    def combinations(self):
        """Every point of the search space as a dict."""
        import itertools
        keys = list(self.search_space)
        return [dict(zip(keys, vals)) for vals in itertools.product(*(self.search_space[k] for k in keys))]
