"""Build system: one nvcc invocation set for sm_100a, one g++ invocation for the host runtime library.

Replaces the reference's multi-backend ``op_builder/`` (``builder.py`` 966 LoC + per-op builders, JIT via
``torch.utils.cpp_extension.load`` with arch-list logic :656-772). There is exactly one target
(``-gencode arch=compute_100a,code=sm_100a``), objects are cached by source mtime, and the outputs land in-tree
(``deepspeed_b200/ops/lib``) so they travel with the source snapshot to GPU machines.
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CSRC = os.path.join(REPO, "csrc")
INCLUDE = os.path.join(REPO, "include")
LIB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
OBJ_DIR = os.path.join(LIB_DIR, "obj")
ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall",
              "--expt-relaxed-constexpr", "-DNDEBUG"]
CXX_FLAGS = ["-O3", "-std=c++17", "-fPIC", "-fopenmp", "-march=native", "-Wall", "-pthread"]


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _cutlass_include():
    """CuTe/CUTLASS header tree vendored in site-packages (used only by kernels that opt in)."""
    for p in sys.path:
        for rel in ("flashinfer/data/cutlass/include", "tilelang/3rdparty/cutlass/include"):
            cand = os.path.join(p, rel)
            if os.path.isdir(cand):
                return cand
    return None


def _stale(src, obj, deps):
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src] + deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout}")
    if verbose and r.stdout.strip():
        print(r.stdout)
    return r.stdout


def cuda_sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def cpu_sources():
    d = os.path.join(CSRC, "cpu")
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(".cpp")) if os.path.isdir(d) else []


def build_cuda_library(verbose=True, force=False, extra_flags=()):
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libdsb200.so``. Returns the library path."""
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = nvcc_path()
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)] if os.path.isdir(INCLUDE) else []
    inc = ["-I", CSRC, "-I", INCLUDE]
    cut = _cutlass_include()
    if cut:
        inc += ["-I", cut]
    jobs = []
    objs = []
    for src in cuda_sources():
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        if force or _stale(src, obj, headers):
            jobs.append([nvcc] + ARCH_FLAGS + NVCC_FLAGS + list(extra_flags) + inc + ["-c", src, "-o", obj])
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda c: _run(c, verbose), jobs))
    out = os.path.join(LIB_DIR, "libdsb200.so")
    if jobs or not os.path.exists(out):
        _run([nvcc, "-shared", "-o", out] + objs + ["-lcudart"] + ARCH_FLAGS, verbose)
    return out


def build_cpu_library(verbose=True, force=False):
    """Host runtime (AVX-512 CPU optimizers, async file I/O thread pool) -> ``libdsb200_cpu.so``."""
    srcs = cpu_sources()
    if not srcs:
        return None
    os.makedirs(LIB_DIR, exist_ok=True)
    out = os.path.join(LIB_DIR, "libdsb200_cpu.so")
    deps = [os.path.join(CSRC, "cpu", f) for f in os.listdir(os.path.join(CSRC, "cpu")) if f.endswith(".h")]
    if force or any(_stale(s, out, deps) for s in srcs):
        _run(["g++"] + CXX_FLAGS + ["-shared", "-o", out] + srcs + ["-I", INCLUDE, "-I", os.path.join(CSRC, "cpu")],
             verbose)
    return out


def build_all(verbose=True, force=False):
    return build_cuda_library(verbose, force), build_cpu_library(verbose, force)


def sass_dump(out_dir, kernels_regex=None):
    """cuobjdump -sass of the built library (evidence of UTC*MMA / UTMALDG etc.) into ``out_dir``."""
    os.makedirs(out_dir, exist_ok=True)
    lib = os.path.join(LIB_DIR, "libdsb200.so")
    txt = subprocess.run(["cuobjdump", "-sass", lib], stdout=subprocess.PIPE, text=True).stdout
    path = os.path.join(out_dir, "libdsb200.sass")
    with open(path, "w") as f:
        f.write(txt)
    return path


if __name__ == "__main__":
    build_all(verbose=True, force="--force" in sys.argv)
