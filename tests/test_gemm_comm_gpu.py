"""csrc/gemm_comm.cu on ONE GPU in loop-back: the "peers" are further buffers on the same device and their flags are
pre-raised, which exercises the per-peer TMA maps, the copy CTAs (bulk async copies + chunk flags), the peer-store
epilogues and the reduce phase without needing a second GPU. The multi-GPU behaviour is tested in
test_symm_multigpu.py."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

AG_A, AG_B, SCATTER_N, RS_M = 1, 2, 4, 8


def _ptrs(*tensors):
    vals = [t.data_ptr() for t in tensors]
    return (ctypes.c_uint64 * 16)(*(vals + [0] * (16 - len(vals))))


class Loop:

    def __init__(self):
        from deepspeed_b200.ops import kernels as K
        assert K.available(), "libdsb200.so must load on a GPU box"
        self.K = K
        self.lib = K.lib()
        P, I, L64, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint
        self.lib.dsb200_gemm_comm_bf16_nt.argtypes = [I, P, P, P, P, P, I, I, I, L64, L64, L64, P, I, I, I, I, I, I, U,
                                                      U, P, P, P, I, I, P]
        self.lib.dsb200_gemm_comm_bf16_nt.restype = I
        self.dev = torch.device("cuda", 0)
        # one flag buffer shared by both "ranks", every channel already far ahead of any epoch we use
        self.flags = torch.full((64 * 32, ), 1 << 20, dtype=torch.int32, device=self.dev)
        self.tickets = torch.zeros(16, dtype=torch.int32, device=self.dev)
        self.chunk_flags = torch.zeros(8192, dtype=torch.int32, device=self.dev)
        self.epoch = 0

    def run(self, mode, a_ptrs, b_ptrs, C, c_ptrs, rs_out, M, N, Kd, lda, ldb, ldc, g_full=None, n_copy=0, rank=0):
        self.epoch += 1
        null = ctypes.c_void_p(0)
        bn = self.lib.dsb200_gemm_comm_bf16_nt(
            mode, a_ptrs, b_ptrs, ctypes.c_void_p(C.data_ptr()) if C is not None else null,
            c_ptrs if c_ptrs is not None else null, ctypes.c_void_p(rs_out.data_ptr()) if rs_out is not None else null,
            M, N, Kd, lda, ldb, ldc, _ptrs(self.flags, self.flags), rank, 2, 9, 10, 11, 12, self.epoch, 0,
            ctypes.c_void_p(self.tickets.data_ptr()), ctypes.c_void_p(g_full.data_ptr()) if g_full is not None else null,
            ctypes.c_void_p(self.chunk_flags.data_ptr()) if g_full is not None else null, n_copy, 0,
            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert bn > 0, self.K.last_error()
        torch.cuda.synchronize()
        return bn


@pytest.fixture(scope="module")
def loop():
    return Loop()


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()


@pytest.mark.parametrize("rank", [0, 1])
@pytest.mark.parametrize("n_copy", [1, 8])
def test_all_gather_b_gemm_loopback(loop, rank, n_copy):
    torch.manual_seed(0)
    M, N, Kd = 384, 1024, 512
    x = torch.randn(M, Kd, device=loop.dev).to(torch.bfloat16)
    W = (torch.randn(N, Kd, device=loop.dev) * 0.05).to(torch.bfloat16)
    shards = [W[:N // 2].contiguous(), W[N // 2:].contiguous()]
    full = torch.zeros(N, Kd, dtype=torch.bfloat16, device=loop.dev)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=loop.dev)
    loop.run(AG_B, _ptrs(x), _ptrs(*shards), out, None, None, M, N, Kd, Kd, Kd, N, g_full=full, n_copy=n_copy,
             rank=rank)
    assert _rel(out, x.float() @ W.float().t()) < 2e-2
    other = 1 - rank
    assert torch.equal(full[other * N // 2:(other + 1) * N // 2], shards[other])  # landed exactly once, bit exact


def test_all_gather_a_gemm_loopback(loop):
    torch.manual_seed(1)
    Mh, N, Kd = 256, 512, 384
    X = torch.randn(2 * Mh, Kd, device=loop.dev).to(torch.bfloat16)
    w = (torch.randn(N, Kd, device=loop.dev) * 0.05).to(torch.bfloat16)
    shards = [X[:Mh].contiguous(), X[Mh:].contiguous()]
    full = torch.zeros(2 * Mh, Kd, dtype=torch.bfloat16, device=loop.dev)
    out = torch.empty(2 * Mh, N, dtype=torch.bfloat16, device=loop.dev)
    loop.run(AG_A, _ptrs(*shards), _ptrs(w), out, None, None, 2 * Mh, N, Kd, Kd, Kd, N, g_full=full, n_copy=4)
    assert _rel(out, X.float() @ w.float().t()) < 2e-2


def test_scatter_and_reduce_scatter_loopback(loop):
    torch.manual_seed(2)
    M, N, Kd = 256, 512, 384
    x = torch.randn(M, Kd, device=loop.dev).to(torch.bfloat16)
    w = (torch.randn(N, Kd, device=loop.dev) * 0.05).to(torch.bfloat16)
    ref = x.float() @ w.float().t()
    # SCATTER_N as rank 0 of 2: column block j lands in buffer j at rows [0, M)
    bufs = [torch.zeros(2 * M, N // 2, dtype=torch.bfloat16, device=loop.dev) for _ in range(2)]
    loop.run(SCATTER_N, _ptrs(x), _ptrs(w), None, _ptrs(*bufs), None, M, N, Kd, Kd, Kd, N // 2)
    for j in range(2):
        assert _rel(bufs[j][:M], ref[:, j * N // 2:(j + 1) * N // 2]) < 2e-2
        assert bufs[j][M:].abs().sum() == 0
    # RS_M as rank 0 of 2: my row block is summed from my staging slots (slot 1 stays zero in loop-back)
    stages = [torch.zeros(2, M // 2, N, dtype=torch.bfloat16, device=loop.dev) for _ in range(2)]
    out = torch.empty(M // 2, N, dtype=torch.bfloat16, device=loop.dev)
    loop.run(RS_M, _ptrs(x), _ptrs(w), None, _ptrs(*stages), out, M, N, Kd, Kd, Kd, N)
    assert _rel(out, ref[:M // 2]) < 2e-2
    assert _rel(stages[1][0], ref[M // 2:]) < 2e-2  # the other rank's row block went to its staging slot 0
