"""NUMA core binding helpers for ``--bind_cores_to_rank`` (reference ``utils/numa.py``)."""
import os
import shutil
import subprocess


def parse_range_list(s):
    """'0-3,8,10-11' -> [0,1,2,3,8,10,11] (sorted, no duplicates)."""
    out = []
    for part in s.split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            a, b = int(a), int(b)
            if b < a:
                raise ValueError(f"Bad range: '{part}', range end must larger than or equal to range start")
            out.extend(range(a, b + 1))
        else:
            out.append(int(part))
    if out != sorted(set(out)):
        raise ValueError(f"Bad range: '{s}', sub ranges must not overlap with each other and should be in ascend order")
    return out


def get_numa_cores():
    try:
        out = subprocess.check_output(["numactl", "--hardware"], text=True)
        nodes = []
        for line in out.splitlines():
            if line.startswith("node") and "cpus:" in line:
                nodes.append([int(c) for c in line.split("cpus:")[1].split()])
        if nodes:
            return nodes
    except Exception:
        pass
    return [list(range(os.cpu_count() or 1))]


def get_numactl_cmd(bind_core_list, num_local_procs, local_rank):
    cores = parse_range_list(bind_core_list) if bind_core_list else list(range(os.cpu_count() or 1))
    per = max(1, len(cores) // num_local_procs)
    mine = cores[local_rank * per:(local_rank + 1) * per]
    if shutil.which("numactl") is None:
        return per, []
    core_str = ",".join(map(str, mine))
    cmd = ["numactl", "-C", core_str]
    for node_id, node_cores in enumerate(get_numa_cores()):
        if set(mine) <= set(node_cores):
            cmd += ["-m", str(node_id)]
            break
    return per, cmd


def parse_range(rng):
    """``"3"`` -> ``range(3, 4)``; ``"2-5"`` -> ``range(2, 6)`` (both ends included)."""
    try:
        v = int(rng)
        return range(v, v + 1)
    except ValueError:
        parts = rng.split("-")
        if len(parts) != 2:
            raise ValueError(f"Bad range: '{rng}', range must be either a number or two number separated by dash")
        lo, hi = int(parts[0]), int(parts[1])
        if lo > hi:
            raise ValueError(f"Bad range: '{rng}', range end must larger than or equal to start")
        return range(lo, hi + 1)


def check_for_numactl_pkg():
    """True when ``numactl`` is on the PATH; otherwise says how to get it and returns False."""
    if shutil.which("numactl") is not None:
        return True
    for pm, hint in (("apt", "apt install numactl"), ("dnf", "dnf install numactl"), ("yum", "yum install numactl")):
        if shutil.which(pm) is not None:
            print(f"numactl is not found; install it with: {hint}")
            break
    else:
        print("numactl is not found on this system")
    return False
