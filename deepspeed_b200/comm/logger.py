"""Comms logger: per-(op, message-size) count / latency / algbw / busbw table.

Reference behaviour: ``deepspeed/utils/comms_logging.py:34`` (calc_bw_log bus-bandwidth formulas per
collective type) and ``:67`` (CommsLogger). Bus-bandwidth conventions follow nccl-tests.
"""
import math


def get_caller_func(frame=3):
    import sys
    return sys._getframe(frame).f_code.co_name


def convert_size(size_bytes):
    if size_bytes == 0:
        return "0B"
    names = ("B", "KB", "MB", "GB", "TB", "PB")
    i = int(math.floor(math.log(size_bytes, 1024)))
    return f"{round(size_bytes / math.pow(1024, i), 2)} {names[i]}"


def calc_bw_log(comm_op, size, duration_ms, n):
    """Return (msg_size_bytes, algbw_Gbps, busbw_Gbps) for one collective call on ``n`` ranks."""
    duration = max(duration_ms, 1e-9) / 1e3
    n = max(n, 1)
    if comm_op in ("all_to_all_single", "all_to_all"):
        tput, busbw = size / duration, (size / duration) * ((n - 1) / n)
    elif comm_op in ("all_gather", "all_gather_into_tensor", "reduce_scatter", "reduce_scatter_tensor"):
        size = size * n if comm_op.startswith("all_gather") and comm_op != "all_gather_into_tensor" else size
        tput, busbw = size / duration, (size / duration) * ((n - 1) / n)
    elif comm_op in ("all_reduce", "all_reduce_coalesced", "inference_all_reduce"):
        tput, busbw = size * 2 / duration, (size / duration) * (2 * (n - 1) / n)
    else:  # send/recv/broadcast/reduce/gather/scatter/barrier
        tput = busbw = size / duration
    return size, tput * 8 / 1e9, busbw * 8 / 1e9


class CommsLogger:
    def __init__(self):
        self.comms_dict = {}
        self.enabled = False
        self.verbose = False
        self.debug = False
        self.prof_all = True
        self.prof_ops = []

    def configure(self, cfg):
        get = (lambda k, d: cfg.get(k, d)) if isinstance(cfg, dict) else (lambda k, d: getattr(cfg, k, d))
        self.enabled = get("enabled", False)
        self.verbose = get("verbose", False)
        self.debug = get("debug", False)
        self.prof_all = get("prof_all", True)
        self.prof_ops = list(get("prof_ops", []) or [])

    def start_profiling_comms(self):
        self.prof_all = True

    def stop_profiling_comms(self):
        self.prof_all = False

    def start_profiling_op(self, ops):
        self.prof_ops = list(set(self.prof_ops) | set(ops))

    def stop_profiling_op(self, ops):
        self.prof_ops = [o for o in self.prof_ops if o not in ops]

    def reset(self):
        self.comms_dict = {}

    def has_data(self):
        return any(per_op for per_op in self.comms_dict.values())

    def get_total_operations(self):
        return sum(ent[0] for per_op in self.comms_dict.values() for ent in per_op.values())

    def get_operation_names(self):
        return [name for name, per_op in self.comms_dict.items() if per_op]

    def append(self, raw_name, record_name, latency_ms, msg_size, world):
        msg_size, algbw, busbw = calc_bw_log(raw_name, msg_size, latency_ms, world)
        per_op = self.comms_dict.setdefault(record_name, {})
        ent = per_op.setdefault(msg_size, [0, [], [], []])
        ent[0] += 1
        ent[1].append(latency_ms)
        ent[2].append(algbw)
        ent[3].append(busbw)
        if self.verbose:
            from ..utils.logging import log_dist
            log_dist(
                f"comm op: {record_name} | time (ms): {latency_ms:.2f} | msg size: {convert_size(msg_size)} | "
                f"algbw (Gbps): {algbw:.2f} | busbw (Gbps): {busbw:.2f}", [0])

    def log_all(self, print_log=True, show_straggler=False, all_reduce_fn=None):
        import torch
        out = {}
        if print_log:
            hdr = f"{'Comm. Op': <20}{'Message Size': <20}{'Count': <10}{'Total Latency(ms)': <20}" \
                  f"{'Avg Latency(ms)': <20}{'tput_avg (Gbps)': <20}{'busbw_avg (Gbps)': <20}"
            print(hdr)
        for name, sizes in self.comms_dict.items():
            if print_log:
                print(name)
            for msg, (count, lats, algs, buss) in sorted(sizes.items()):
                tot, avg = sum(lats), sum(lats) / len(lats)
                alg, bus = sum(algs) / len(algs), sum(buss) / len(buss)
                out.setdefault(name, {})[msg] = dict(count=count, total_ms=tot, avg_ms=avg, algbw=alg, busbw=bus)
                if print_log:
                    print(f"{' ': <20}{convert_size(msg): <20}{count: <10}{tot: <20.2f}{avg: <20.2f}"
                          f"{alg: <20.2f}{bus: <20.2f}")
        if show_straggler and all_reduce_fn is not None:
            if print_log:
                print("\nStraggler effect (total latency - min-over-ranks latency):")
            for name, sizes in self.comms_dict.items():
                for msg, (count, lats, _, _) in sorted(sizes.items()):
                    t = torch.tensor(lats, dtype=torch.float64)
                    if torch.cuda.is_available():
                        t = t.cuda()
                    mn = t.clone()
                    import torch.distributed as dist
                    all_reduce_fn(mn, op=dist.ReduceOp.MIN)
                    strag = float((t - mn).sum())
                    out[name][msg]["straggler_ms"] = strag
                    if print_log:
                        print(f"{name: <20}{convert_size(msg): <20}{count: <10}{strag: <20.2f}")
        return out
