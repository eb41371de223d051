"""dsb200_sumsq_noalloc through the C ABI.

The entry point exists for callers that cannot give the library scratch memory: a process whose memory pool already
owns all of the GPU's HBM (a raw cudaMalloc would fail or make the pool shrink), a latency-critical loop that must
not hit the device-wide synchronisation of cudaMalloc/cudaFree, or a bare C program with no allocator to pass
along. It keeps its partial sums in a small __device__ array inside the library. This test loads the shared object
with ctypes (no Python wrapper in between), calls the new entry point twice in a row and checks both results
against the existing dsb200_sumsq, which gets a caller-provided scratch buffer. torch only supplies the input and
output buffers here; the entry point itself allocates nothing.

The file name sorts last on purpose: a run that stops at the first failure has then already run everything else.
"""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "deepspeed_b200", "ops", "lib",
                   "libdsb200.so")
MAX_ELEMS = 1 << 34
DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _lib():
    L = ctypes.CDLL(LIB)
    P, I, L64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
    L.dsb200_sumsq_noalloc.argtypes = [P, L64, I, P, P]
    L.dsb200_sumsq_noalloc.restype = I
    L.dsb200_sumsq.argtypes = [P, L64, I, P, P, P]
    L.dsb200_sumsq.restype = I
    L.dsb200_last_error.restype = ctypes.c_char_p
    return L


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(dtype):
    L = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(0)
    scratch = torch.zeros(2048, dtype=torch.float32, device=dev)  # for the existing entry point only
    for call, n in enumerate((3_000_017, 70_001)):  # second call reuses the array the first one left behind
        x = torch.randn(n, device=dev).to(dtype)
        got = torch.full((1, ), -1.0, dtype=torch.float32, device=dev)
        ref = torch.full((1, ), -2.0, dtype=torch.float32, device=dev)
        rc = L.dsb200_sumsq_noalloc(_ptr(x), n, DT[dtype], _ptr(got), stream)
        assert rc == 0, L.dsb200_last_error()
        rc = L.dsb200_sumsq(_ptr(x), n, DT[dtype], _ptr(scratch), _ptr(ref), stream)
        assert rc == 0, L.dsb200_last_error()
        torch.cuda.synchronize()
        exact = float(x.double().pow(2).sum())
        assert abs(float(got) - float(ref)) <= 2e-4 * abs(float(ref)) + 1e-6, (call, float(got), float(ref))
        assert abs(float(got) - exact) <= 2e-4 * exact + 1e-6, (call, float(got), exact)


def test_noalloc_rejects_oversize_and_bad_dtype():
    L = _lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = torch.zeros(1, dtype=torch.float32, device="cuda")
    rc = L.dsb200_sumsq_noalloc(ctypes.c_void_p(0), MAX_ELEMS + 1, 2, _ptr(out), stream)
    assert rc != 0 and b"2^34" in L.dsb200_last_error()
    rc = L.dsb200_sumsq_noalloc(ctypes.c_void_p(0), 16, 7, _ptr(out), stream)
    assert rc != 0 and b"dtype" in L.dsb200_last_error()
    x = torch.ones(0, device="cuda")
    rc = L.dsb200_sumsq_noalloc(_ptr(out), 0, 0, _ptr(out), stream)  # empty input -> 0, no launch
    torch.cuda.synchronize()
    assert rc == 0 and float(out) == 0.0
