"""Symmetric-memory layer: peer-mapped buffers, NVLS multicast, epoch flags, and the fused collective launchers.

Reference analogue: ``comm/torch.py:455 enable_symm_mem_for_group`` + ``csrc/compile/z3.cpp:95-113`` (barrier ->
per-peer ``copy_`` -> barrier). Here a :class:`SymmContext` (one per process group) owns

* a symmetric *flag* buffer (epoch counters written with ``st.release.sys`` by peers, see ``csrc/symm.cu``),
* a pool of symmetric data buffers allocated through ``torch.distributed._symmetric_memory`` (rendezvous gives the
  peer pointer table and, when the fabric supports it, the NVSwitch multicast pointer),
* per-channel monotonically increasing epochs (every rank issues the same sequence of collectives per channel).

and the launchers :meth:`reduce_scatter`, :meth:`all_gather`, :meth:`all_to_all` run ONE kernel that moves the
data over NVLink from inside the kernel and fuses the adjacent elementwise work.
Availability: CUDA, single NVSwitch domain, world <= 16, ``torch.distributed._symmetric_memory`` usable. Otherwise
callers use the NCCL path in :mod:`deepspeed_b200.comm.comm` (also the numerical oracle in tests).
"""
import ctypes
import os

import torch

from ..ops import kernels as K
from ..utils.logging import log_dist, logger
from . import comm as dist

CH_RS_READY, CH_RS_DONE, CH_AG_READY, CH_AG_DONE, CH_A2A_READY, CH_A2A_DONE, CH_USER = 0, 1, 2, 3, 4, 5, 6
CH_QAG_READY, CH_QAG_DONE, CH_GIN_READY, CH_GIN_DONE = 7, 8, 9, 10
CH_GA2A_READY, CH_GA2A_DONE, CH_GRS_READY, CH_GRS_DONE = 11, 12, 13, 14
CH_CE_READY, CH_CE_DONE = 15, 16
GEMM_AG_A, GEMM_AG_B, GEMM_SCATTER_N, GEMM_RS_M = 1, 2, 4, 8
_CONTEXTS = {}


def _declare(L):
    P, I, L64, F, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_uint
    U64 = ctypes.c_uint64
    L.dsb200_symm_reduce_scatter.argtypes = [P, P, U64, I, I, L64, I, P, I, I, F, P, P, I, I, U, I, P]
    L.dsb200_symm_reduce_scatter.restype = I
    L.dsb200_symm_all_gather.argtypes = [P, P, U64, I, I, L64, P, P, P, I, I, U, U, I, P]
    L.dsb200_symm_all_gather.restype = I
    L.dsb200_symm_all_to_all.argtypes = [P, P, I, I, L64, L64, I, P, P, I, I, U, I, P]
    L.dsb200_symm_all_to_all.restype = I
    L.dsb200_symm_rs_adam.argtypes = [P, P, I, I, L64, I, F, P, P, P, P, I, P, P, P, P, I, I, U, I, P]
    L.dsb200_symm_rs_adam.restype = I
    L.dsb200_symm_fp8_stage_bytes.argtypes = [L64, I]
    L.dsb200_symm_fp8_stage_bytes.restype = L64
    L.dsb200_symm_all_gather_fp8.argtypes = [P, I, P, P, I, I, L64, I, P, P, I, I, U, U, I, P]
    L.dsb200_symm_all_gather_fp8.restype = I
    L.dsb200_gemm_comm_bf16_nt.argtypes = [I, P, P, P, P, P, I, I, I, L64, L64, L64, P, I, I, I, I, I, I, U, U, P, P, P, I,
                                           I, P]
    L.dsb200_gemm_comm_bf16_nt.restype = I
    L.dsb200_symm_all_gather_ce.argtypes = [P, I, I, L64, P, P, I, P]
    L.dsb200_symm_all_gather_ce.restype = I
    L.dsb200_symm_wait.argtypes = [P, I, I, I, U, P]
    L.dsb200_symm_wait.restype = I
    L.dsb200_symm_signal.argtypes = [P, I, I, I, U, P]
    L.dsb200_symm_signal.restype = I


def is_supported(group=None):
    """Can the fused NVLink collectives be used for this group?"""
    if os.environ.get("DSB200_DISABLE_SYMM", "0") == "1":
        return False
    if not (torch.cuda.is_available() and dist.is_initialized()):
        return False
    world = dist.get_world_size(group)
    if world < 2 or world > 16 or dist.get_backend(group) != "nccl":
        return False
    try:
        import importlib
        importlib.import_module("torch.distributed._symmetric_memory")
    except Exception:
        return False
    return K.available()


class SymmBuffer:
    """A symmetric allocation: same size on every rank, every rank can address all copies."""

    def __init__(self, ctx, nbytes, want_multicast=True):
        import torch.distributed._symmetric_memory as symm_mem
        self.ctx = ctx
        self.nbytes = int((nbytes + 255) // 256 * 256)
        self.tensor = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=ctx.device)
        self.handle = symm_mem.rendezvous(self.tensor, ctx.group)
        self.ptrs = (ctypes.c_uint64 * 16)(*([int(p) for p in self.handle.buffer_ptrs] + [0] * (16 - ctx.world)))
        mc = 0
        if want_multicast:
            try:
                mc = int(self.handle.multicast_ptr or 0) if self.handle.has_multicast_support else 0
            except Exception:
                mc = 0
        self.mc_ptr = mc
        self.last_rs_epoch = 0  # epoch of the last reduce-scatter that read this buffer

    def view(self, numel, dtype, offset_bytes=0):
        esz = torch.empty((), dtype=dtype).element_size()
        return self.tensor[offset_bytes:offset_bytes + numel * esz].view(dtype)


class SymmContext:

    def __init__(self, group):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = K.lib()
        _declare(self.lib)
        self.flags = SymmBuffer(self, 64 * 32 * 4, want_multicast=False)
        self.flags.tensor.zero_()
        self.tickets = torch.zeros(64, dtype=torch.int32, device=self.device)
        self.partials = torch.zeros(64 * 16, dtype=torch.float32, device=self.device)
        self.epoch = [0] * 64
        # CTAs per communication kernel: enough loads in flight to fill NVLink (~2 us x 900 GB/s = 1.8 MB ->
        # 16-32 CTAs x 512 threads x 16 B x unroll), few enough to leave the SMs to the overlapped GEMMs
        self.default_ctas = int(os.environ.get("DSB200_COMM_CTAS", "24"))
        torch.cuda.synchronize()
        dist.barrier(group=group)
        self.multicast = False
        log_dist(f"SymmContext: world={self.world} symmetric memory ready", [0])

    @staticmethod
    def get(group):
        key = id(group)
        if key not in _CONTEXTS:
            _CONTEXTS[key] = SymmContext(group)
        return _CONTEXTS[key]

    def alloc(self, nbytes, want_multicast=True):
        return SymmBuffer(self, nbytes, want_multicast)

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _next(self, ch):
        self.epoch[ch] += 1
        return self.epoch[ch]

    # ---- collectives ---------------------------------------------------------------------------
    def reduce_scatter(self, buf: SymmBuffer, in_dtype, shard_numel, out, accumulate, scale, use_multicast=None,
                       ctas=None, want_sumsq=False):
        """out (+)= scale * sum_p buf_p[rank*shard : (rank+1)*shard]. Returns epoch (for DONE waits)."""
        ep = self._next(CH_RS_READY)
        mc = buf.mc_ptr if (use_multicast if use_multicast is not None else True) else 0
        self._presync_rs(ep)
        rc = self.lib.dsb200_symm_reduce_scatter(
            buf.ptrs, self.flags.ptrs, ctypes.c_uint64(mc if in_dtype == torch.bfloat16 else 0), self.rank, self.world,
            shard_numel, K._DT[in_dtype], ctypes.c_void_p(out.data_ptr()), K._DT[out.dtype], 1 if accumulate else 0,
            float(scale), ctypes.c_void_p(self.partials.data_ptr()) if want_sumsq else ctypes.c_void_p(0),
            ctypes.c_void_p(self.tickets.data_ptr()), CH_RS_READY, CH_RS_DONE, ep,
            self.default_ctas if ctas is None else ctas, self._stream())
        if rc < 0:
            raise RuntimeError(f"symm reduce_scatter failed: {K.last_error()}")
        K._LAUNCHES += 1
        buf.last_rs_epoch = ep
        return ep

    def _presync_rs(self, ep):
        """Wait for the peers' buckets with two one-warp kernels BEFORE the multi-CTA reduce-scatter kernel starts: the
        kernel's own READY handshake is then already satisfied, so its CTAs hold SM slots only while they move data
        (profile at 8 GPUs: 1.6 ms per launch, almost all of it spinning for the slowest rank next to the GEMMs)."""
        if os.environ.get("DSB200_RS_PRESYNC", "1") == "1":
            self._signal(CH_RS_READY, ep)
            self._wait(CH_RS_READY, ep)

    def rs_adam(self, buf: SymmBuffer, in_dtype, shard_numel, grad_scale, p, m, v, lp, seg_bounds, hyper,
                want_sumsq=False, ctas=None):
        """Fused path (1): reduce-scatter my gradient shard out of every peer's bucket ``buf`` and apply AdamW to the
        fp32 master ``p`` / moments ``m``, ``v`` (+ bf16 copy ``lp``) in the same kernel. ``seg_bounds`` (len nseg+1,
        multiples of 8) and ``hyper`` (nseg rows of lr, b1, b2, eps, wd, bc1, bc2, mode) describe the parameter-group
        segments of the shard. Returns (epoch, grid)."""
        ep = self._next(CH_RS_READY)
        self._presync_rs(ep)
        nseg = len(seg_bounds) - 1
        bounds = (ctypes.c_int64 * (nseg + 1))(*[int(b) for b in seg_bounds])
        hy = (ctypes.c_float * (nseg * 8))(*[float(x) for row in hyper for x in row])
        grid = self.lib.dsb200_symm_rs_adam(
            buf.ptrs, self.flags.ptrs, self.rank, self.world, shard_numel, K._DT[in_dtype], float(grad_scale),
            ctypes.c_void_p(p.data_ptr()), ctypes.c_void_p(m.data_ptr()), ctypes.c_void_p(v.data_ptr()),
            ctypes.c_void_p(lp.data_ptr() if lp is not None else 0), nseg, bounds, hy,
            ctypes.c_void_p(self.partials.data_ptr()) if want_sumsq else ctypes.c_void_p(0),
            ctypes.c_void_p(self.tickets.data_ptr()), CH_RS_READY, CH_RS_DONE, ep,
            self.default_ctas if ctas is None else ctas, self._stream())
        if grid < 0:
            raise RuntimeError(f"symm rs_adam failed: {K.last_error()}")
        K._LAUNCHES += 1
        buf.last_rs_epoch = ep
        return ep, grid

    def fp8_stage_bytes(self, shard_elems, block=256):
        return int(self.lib.dsb200_symm_fp8_stage_bytes(shard_elems, block))

    def all_gather_fp8(self, my_shard, stage: SymmBuffer, full, block=256, ctas=None):
        """Fused path (3), qwZ: quantize ``my_shard`` (bf16/fp16) to block-scaled e4m3 in ``stage``, pull the peers'
        payload + scales over NVLink and dequantize into ``full`` -- one kernel, half the NVLink bytes. Every rank
        (the owner included) ends up with the same dequantized values."""
        ep = self._next(CH_QAG_READY)
        n = my_shard.numel()
        assert stage.nbytes >= self.fp8_stage_bytes(n, block) and full.numel() >= n * self.world
        rc = self.lib.dsb200_symm_all_gather_fp8(
            ctypes.c_void_p(my_shard.data_ptr()), K._DT[my_shard.dtype], stage.ptrs, self.flags.ptrs, self.rank,
            self.world, n, block, ctypes.c_void_p(full.data_ptr()), ctypes.c_void_p(self.tickets.data_ptr() + 12),
            CH_QAG_READY, CH_QAG_DONE, ep, getattr(stage, "last_qag_epoch", 0),
            self.default_ctas if ctas is None else ctas, self._stream())
        if rc != 0:
            raise RuntimeError(f"symm all_gather_fp8 failed: {K.last_error()}")
        K._LAUNCHES += 1
        stage.last_qag_epoch = ep
        return ep

    # ---- tcgen05 GEMM fused with a collective (csrc/gemm_comm.cu) -------------------------------------
    def _signal(self, channel, epoch):
        self.lib.dsb200_symm_signal(self.flags.ptrs, self.rank, self.world, channel, epoch, self._stream())
        K._LAUNCHES += 1

    def _wait(self, channel, epoch):
        if epoch > 0:
            self.lib.dsb200_symm_wait(self.flags.ptrs, self.rank, self.world, channel, epoch, self._stream())
            K._LAUNCHES += 1

    def _gemm_comm(self, mode, a_ptrs, b_ptrs, out, c_ptrs, rs_out, M, N, Kd, lda, ldb, ldc, chans, epoch, reuse,
                   max_ctas=0, g_full=None, n_copy=0):
        null = ctypes.c_void_p(0)
        if g_full is not None and getattr(self, "chunk_flags", None) is None:
            self.chunk_flags = torch.zeros(8192, dtype=torch.int32, device=self.device)
        bn = self.lib.dsb200_gemm_comm_bf16_nt(
            mode, a_ptrs, b_ptrs, ctypes.c_void_p(out.data_ptr()) if out is not None else null,
            c_ptrs if c_ptrs is not None else null, ctypes.c_void_p(rs_out.data_ptr()) if rs_out is not None else null,
            M, N, Kd, lda, ldb, ldc, self.flags.ptrs, self.rank, self.world, chans[0], chans[1], chans[2], chans[3],
            epoch, reuse, ctypes.c_void_p(self.tickets.data_ptr() + 32),
            ctypes.c_void_p(g_full.data_ptr()) if g_full is not None else null,
            ctypes.c_void_p(self.chunk_flags.data_ptr()) if g_full is not None else null, n_copy, max_ctas,
            self._stream())
        if bn < 0:
            raise RuntimeError(f"gemm_comm failed: {K.last_error()}")
        K._LAUNCHES += 1
        return bn

    @staticmethod
    def _one(t):
        return (ctypes.c_uint64 * 16)(*([t.data_ptr()] + [0] * 15))

    def all_gather_gemm(self, x, w_buf: SymmBuffer, n_total, out=None, gathered="b", max_ctas=0, full=None,
                        n_copy=None):
        """Fused path 2. ``gathered="b"``: ``out[M, n_total] = x[M, K] @ W^T`` where rank p holds rows
        ``[p*n_total/world, ...)`` of ``W`` in ``w_buf`` - the ZeRO-3 weight all-gather happens tile by tile inside the
        GEMM. ``gathered="a"``: ``out[m_total, N] = X @ w^T`` with the rows of ``X`` spread over the ranks' ``w_buf``
        (sequence-parallel input of an AutoTP column-parallel linear); then ``x`` is the local weight [N, K]."""
        Kd = x.shape[1]
        ep = self._next(CH_GIN_READY)
        self._signal(CH_GIN_READY, ep)  # stream order: my shard is final before peers may read it
        if gathered == "b":
            M, N = x.shape[0], n_total
            a_ptrs, b_ptrs, mode = self._one(x), w_buf.ptrs, GEMM_AG_B
        else:
            M, N = n_total, x.shape[0]
            a_ptrs, b_ptrs, mode = w_buf.ptrs, self._one(x), GEMM_AG_A
        if out is None:
            out = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
        if full is None:  # local landing buffer of the gathered operand (for ZeRO-3: the unit's full parameter)
            cache = self.__dict__.setdefault("_gather_full", {})
            full = cache.get(n_total * Kd)
            if full is None:
                full = cache[n_total * Kd] = torch.empty(n_total * Kd, dtype=torch.bfloat16, device=x.device)
        if n_copy is None:
            n_copy = int(os.environ.get("DSB200_GEMM_COPY_CTAS", "12"))
        self._gemm_comm(mode, a_ptrs, b_ptrs, out, None, None, M, N, Kd, Kd if gathered == "a" else x.stride(0),
                        Kd if gathered == "b" else x.stride(0), out.stride(0),
                        (CH_GIN_READY, CH_GIN_DONE, 0, 0), ep, 0, max_ctas, g_full=full, n_copy=n_copy)
        w_buf.last_gin_epoch = ep
        return out

    def wait_gemm_readers(self, buf: SymmBuffer):
        """Before overwriting a shard that fused GEMMs read: wait until every peer's kernel finished reading it."""
        self._wait(CH_GIN_DONE, getattr(buf, "last_gin_epoch", 0))

    def gemm_all_to_all(self, x, w, out_buf: SymmBuffer, max_ctas=0):
        """Fused path 4. ``Y = x[M, K] @ w[N, K]^T`` with column block j of ``Y`` delivered to rank j: afterwards
        ``out_buf`` on every rank holds ``[world * M, N / world]`` (rows ordered by source rank) - the Ulysses
        sequence-sharded -> head-sharded re-layout done by the GEMM epilogue. Call :meth:`release_a2a_out` once the
        consumer has read ``out_buf``."""
        M, Kd = x.shape
        N = w.shape[0]
        ep = self._next(CH_GA2A_READY)
        self._gemm_comm(GEMM_SCATTER_N, self._one(x), self._one(w), None, out_buf.ptrs, None, M, N, Kd, x.stride(0),
                        w.stride(0), N // self.world, (0, 0, CH_GA2A_READY, CH_GA2A_DONE), ep,
                        getattr(out_buf, "last_ga2a_done", 0), max_ctas)
        out_buf.last_ga2a_epoch = ep
        return out_buf.view(self.world * M * (N // self.world), torch.bfloat16).view(self.world * M, N // self.world)

    def release_a2a_out(self, out_buf: SymmBuffer):
        ep = getattr(out_buf, "last_ga2a_epoch", 0)
        if ep:
            self._signal(CH_GA2A_DONE, ep)
            out_buf.last_ga2a_done = ep

    def gemm_reduce_scatter(self, x, w, stage_buf: SymmBuffer, out=None, max_ctas=0):
        """Fused AutoTP row-parallel linear: every rank computes the partial ``x[M, K_local] @ w[N, K_local]^T``;
        row block j goes to rank j's staging slot while the GEMM runs, then each rank sums its ``world`` slots:
        ``out[M / world, N] = sum_p partial_p[rank rows]``. ``stage_buf`` needs ``M * N * 2`` bytes."""
        M, Kd = x.shape
        N = w.shape[0]
        if out is None:
            out = torch.empty((M // self.world, N), dtype=torch.bfloat16, device=x.device)
        ep = self._next(CH_GRS_READY)
        self._gemm_comm(GEMM_RS_M, self._one(x), self._one(w), None, stage_buf.ptrs, out, M, N, Kd, x.stride(0),
                        w.stride(0), out.stride(0), (0, 0, CH_GRS_READY, CH_GRS_DONE), ep,
                        getattr(stage_buf, "last_grs_epoch", 0), max_ctas)
        stage_buf.last_grs_epoch = ep
        return out

    def wait_rs_done(self, epoch):
        """Enqueue (current stream) a wait until every peer finished reading my buffer in RS ``epoch``."""
        if epoch > 0:
            self.lib.dsb200_symm_wait(self.flags.ptrs, self.rank, self.world, CH_RS_DONE, epoch, self._stream())
            K._LAUNCHES += 1

    def all_gather(self, shard_buf: SymmBuffer, shard_offset_bytes, shard_bytes, full, ctas=None):
        """full[p*shard_bytes ...] = peer p's bytes at ``shard_offset_bytes`` of ``shard_buf`` (pull over NVLink)."""
        ep = self._next(CH_AG_READY)
        ptrs = (ctypes.c_uint64 * 16)(*[(int(p) + shard_offset_bytes) if p else 0 for p in shard_buf.ptrs])
        rc = self.lib.dsb200_symm_all_gather(ptrs, self.flags.ptrs, ctypes.c_uint64(0), self.rank, self.world,
                                             shard_bytes, ctypes.c_void_p(0), ctypes.c_void_p(full.data_ptr()),
                                             ctypes.c_void_p(self.tickets.data_ptr() + 4), CH_AG_READY, CH_AG_DONE, ep,
                                             0, self.default_ctas if ctas is None else ctas, self._stream())
        if rc != 0:
            raise RuntimeError(f"symm all_gather failed: {K.last_error()}")
        K._LAUNCHES += 1
        return ep

    # ---- copy-engine all-gather (ZeRO-3 parameter gather) ------------------------------------------------
    # The parameter shards live in symmetric memory and only change at the optimizer step, so the gather needs no
    # per-call handshake: the writer publishes once (`publish_shards`), the first gather afterwards waits for every
    # peer's publication, and the copies themselves are plain DMA pulls over NVLink that occupy no SM -- the GEMMs they
    # overlap with keep the whole GPU. Before the shards are rewritten the writer waits until every peer has finished
    # the gathers issued so far (`wait_ce_readers`; all ranks issue the same sequence of gathers).
    def _ce_state(self):
        if not hasattr(self, "_ce_helpers"):
            n = int(os.environ.get("DSB200_CE_STREAMS", str(min(self.world - 1, 7))))
            self._ce_helpers = [torch.cuda.Stream(priority=-1) for _ in range(max(0, min(n, 8)))]
            self._ce_helper_arr = (ctypes.c_void_p * 8)(*([s.cuda_stream for s in self._ce_helpers] +
                                                          [None] * (8 - len(self._ce_helpers))))
            self._ce_pub, self._ce_seen, self._ce_issued = 0, 0, 0
        return self

    def publish_shards(self):
        """Call on the stream that (re)wrote symmetric parameter shards, after the writes."""
        self._ce_state()
        self._ce_pub = self._next(CH_CE_READY)
        self._signal(CH_CE_READY, self._ce_pub)

    def wait_ce_readers(self):
        """Call on the stream that is about to rewrite symmetric parameter shards."""
        self._ce_state()
        if self._ce_issued:
            self._wait(CH_CE_DONE, self._ce_issued)

    def all_gather_ce(self, shard_buf: SymmBuffer, shard_offset_bytes, shard_bytes, full, need_epoch=0):
        """full[p*shard_bytes ...] = peer p's bytes at ``shard_offset_bytes`` of ``shard_buf``, on the copy engines.
        ``need_epoch``: publication after which this shard is final on every peer (a pipelined optimizer step
        publishes unit by unit); 0 = the latest publication."""
        self._ce_state()
        need = need_epoch if 0 < need_epoch <= self._ce_pub else self._ce_pub
        if need > self._ce_seen:
            self._wait(CH_CE_READY, need)  # every peer's copy of this shard is final
            self._ce_seen = need
        ptrs = (ctypes.c_uint64 * 16)(*[(int(p) + shard_offset_bytes) if p else 0 for p in shard_buf.ptrs])
        rc = self.lib.dsb200_symm_all_gather_ce(ptrs, self.rank, self.world, shard_bytes,
                                                ctypes.c_void_p(full.data_ptr()), self._ce_helper_arr,
                                                len(self._ce_helpers), self._stream())
        if rc != 0:
            raise RuntimeError(f"symm all_gather_ce failed: {K.last_error()}")
        self._ce_issued += 1
        self._signal(CH_CE_DONE, self._ce_issued)

    def all_to_all(self, in_buf: SymmBuffer, rows, row_bytes, out, interleave=False, ctas=None):
        """out[p] = peer p's chunk ``rank`` of its input buffer viewed as [world, rows, row_bytes]."""
        ep = self._next(CH_A2A_READY)
        rc = self.lib.dsb200_symm_all_to_all(in_buf.ptrs, self.flags.ptrs, self.rank, self.world, rows, row_bytes,
                                             1 if interleave else 0, ctypes.c_void_p(out.data_ptr()),
                                             ctypes.c_void_p(self.tickets.data_ptr() + 8), CH_A2A_READY, CH_A2A_DONE,
                                             ep, self.default_ctas if ctas is None else ctas, self._stream())
        if rc != 0:
            raise RuntimeError(f"symm all_to_all failed: {K.last_error()}")
        K._LAUNCHES += 1
        in_buf.last_a2a_epoch = ep
        return ep

    def wait_a2a_done(self, epoch):
        if epoch > 0:
            self.lib.dsb200_symm_wait(self.flags.ptrs, self.rank, self.world, CH_A2A_DONE, epoch, self._stream())
            K._LAUNCHES += 1

    def wait_ag_done(self, epoch):
        if epoch > 0:
            self.lib.dsb200_symm_wait(self.flags.ptrs, self.rank, self.world, CH_AG_DONE, epoch, self._stream())
            K._LAUNCHES += 1


class SymmSlotPool:
    """Pool of symmetric buffers (gradient buckets / a2a staging) with ownership tracking. A slot handed out by
    ``acquire`` stays busy until ``release``; before a released slot is rewritten the producer stream waits for the
    peers' DONE flag of the collective that last read it. ``acquire`` returns None when every slot is held (a unit
    whose gradients are still incomplete keeps its slot until the backward epilogue): the caller then uses a private
    buffer and the NCCL path. Every rank sees the same hook order, so every rank takes the same decision."""

    def __init__(self, ctx: SymmContext, slot_bytes, n_slots=3, kind="rs"):
        self.ctx, self.kind = ctx, kind
        self.slots = [ctx.alloc(slot_bytes) for _ in range(n_slots)]
        self.busy = [False] * n_slots
        self.next = 0
        self.slot_bytes = self.slots[0].nbytes

    def acquire(self):
        """Next free slot in round-robin order (None if all are held); enqueues the DONE wait of its previous use
        on the current stream."""
        n = len(self.slots)
        for k in range(n):
            i = (self.next + k) % n
            if not self.busy[i]:
                break
        else:
            return None
        self.busy[i] = True
        self.next = (i + 1) % n
        s = self.slots[i]
        if self.kind == "rs":
            self.ctx.wait_rs_done(getattr(s, "last_rs_epoch", 0))
        else:
            self.ctx.wait_a2a_done(getattr(s, "last_a2a_epoch", 0))
        return s

    def release(self, slot):
        """The holder has enqueued the collective that consumes ``slot`` (or dropped it); it may be handed out again."""
        for i, s in enumerate(self.slots):
            if s is slot:
                self.busy[i] = False
                return

    def n_free(self):
        return self.busy.count(False)
