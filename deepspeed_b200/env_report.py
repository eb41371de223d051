"""``ds_report``: environment + native-op status (reference ``env_report.py``)."""
import os
import shutil
import subprocess
import sys

GREEN, RED, YELLOW, END = "\033[92m", "\033[91m", "\033[93m", "\033[0m"
OKAY, NO = f"{GREEN}[OKAY]{END}", f"{RED}[NO]{END}"


def op_report(verbose=True):
    from .ops import builder
    rows = []
    lib = os.path.join(builder.LIB_DIR, "libdsb200.so")
    cpu = os.path.join(builder.LIB_DIR, "libdsb200_cpu.so")
    for src in builder.cuda_sources():
        name = os.path.basename(src)[:-3]
        obj = os.path.join(builder.OBJ_DIR, name + ".o")
        rows.append((name, os.path.exists(obj), "sm_100a CUDA"))
    for src in builder.cpu_sources():
        rows.append((os.path.basename(src)[:-4], os.path.exists(cpu), "host C++"))
    print("-" * 70 + "\nnative op report (built objects in deepspeed_b200/ops/lib)\n" + "-" * 70)
    print(f"{'op name':<24}{'built':<12}{'kind'}")
    for name, ok, kind in rows:
        print(f"{name:<24}{(OKAY if ok else NO):<21}{kind}")
    print(f"{'libdsb200.so':<24}{(OKAY if os.path.exists(lib) else NO)}")
    print(f"{'libdsb200_cpu.so':<24}{(OKAY if os.path.exists(cpu) else NO)}")
    return rows


def nvcc_version():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    try:
        out = subprocess.check_output([nvcc, "-V"], text=True)
        return out.strip().split("release")[-1].split(",")[0].strip()
    except Exception:
        return "not found"


def debug_report():
    import torch

    from .version import __reference_api__, __version__
    rows = [("torch install path", torch.__path__), ("torch version", torch.__version__),
            ("deepspeed_b200 install path", [os.path.dirname(os.path.abspath(__file__))]),
            ("deepspeed_b200 info", f"{__version__} (reference API {__reference_api__})"),
            ("torch cuda version", torch.version.cuda), ("nvcc version", nvcc_version()),
            ("cuda available", torch.cuda.is_available()),
            ("device", torch.cuda.get_device_name(0) if torch.cuda.is_available() else "cpu"),
            ("nccl", ".".join(map(str, torch.cuda.nccl.version())) if torch.cuda.is_available() else "n/a"),
            ("python", sys.version.split()[0])]
    shm, warns = get_shm_size()
    rows.append(("shared memory (/dev/shm) size", shm))
    print("DeepSpeed-B200 general environment info:")
    for name, value in rows:
        print(f"{name} {'.' * max(1, 34 - len(name))} {value}")
    for w in warns or []:
        print(w)


def human_readable_size(size):
    units = ["B", "KB", "MB", "GB", "TB"]
    i = 0
    while size >= 1024 and i < len(units) - 1:
        size /= 1024
        i += 1
    return f"{size:.2f} {units[i]}"


def get_shm_size():
    """(human readable size of /dev/shm, list of warnings): NCCL and the shared-memory all-reduce need it to be large
    enough (docker defaults to 64 MB)."""
    try:
        st = os.statvfs("/dev/shm")
    except (OSError, ValueError, AttributeError):
        return "UNKNOWN", None
    size = st.f_frsize * st.f_blocks
    warn = []
    if size < 512 * 1024**2:
        warn.append(" [WARNING] /dev/shm size might be too small, if running in docker increase to at least "
                    "--shm-size='1gb'")
        warn.append(" [WARNING] see more details about NCCL requirements: "
                    "https://docs.nvidia.com/deeplearning/nccl/user-guide/docs/troubleshooting.html#sharing-data")
    return human_readable_size(size), warn


def parse_arguments():
    import argparse
    p = argparse.ArgumentParser()
    p.add_argument("--hide_operator_status", action="store_true",
                   help="Suppress display of installation and compatibility statuses of the operators.")
    p.add_argument("--hide_errors_and_warnings", action="store_true", help="Suppress warning and error messages.")
    return p.parse_args()


def main(hide_operator_status=False, hide_errors_and_warnings=False):
    if not hide_operator_status:
        op_report(verbose=not hide_errors_and_warnings)
    debug_report()


def cli_main():
    a = parse_arguments()
    main(a.hide_operator_status, a.hide_errors_and_warnings)


if __name__ == "__main__":
    main()
