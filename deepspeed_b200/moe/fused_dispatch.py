"""Expert-parallel dispatch / combine on the fused NVLink kernels of ``csrc/moe_comm.cu``.

The reference moves routed tokens with ``index_select`` -> ``all_to_all_single`` (variable splits) -> permute, runs the
experts, and returns them with un-permute -> ``all_to_all_single`` -> scale -> ``index_add``
(``module_inject/auto_ep_layer.py:184-350,575-730``). Here each direction is ONE kernel that reads peer memory:

* ``dispatch``: the expert rank pulls the rows routed to its experts out of every peer's activation buffer, indexed by
  that peer's expert-sorted token list, and writes them expert-contiguous;
* ``combine``: the token owner pulls the ``k`` expert outputs of each token from the expert ranks, applies the router
  weights and writes the token row once.

Each is the other's backward (combine with unit weights is the backward of dispatch; dispatch with the router weights as
per-slot scale is the backward of combine), so training needs no further kernels. One host sync per layer remains (the
``[ep, E]`` count matrix sizes the expert buffers -- the same sync the reference has, ``auto_ep_layer.py:95-133``).
"""
import ctypes
from dataclasses import dataclass
from typing import List

import torch

from .. import comm as dist
from ..comm import symm as _symm
from ..ops import kernels as K

CH_D_READY, CH_D_DONE, CH_C_READY, CH_C_DONE = 17, 18, 19, 20
_DECLARED = False
_COMMS = {}


def _lib():
    global _DECLARED
    L = K.lib()
    if not _DECLARED:
        P, I, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_uint
        L.dsb200_moe_dispatch.argtypes = [P, P, P, P, I, I, P, I, I, I, P, P, I, I, U, I, P]
        L.dsb200_moe_dispatch.restype = I
        L.dsb200_moe_combine.argtypes = [P, P, I, I, P, P, P, I, I, I, P, P, P, I, I, U, I, P]
        L.dsb200_moe_combine.restype = I
        _DECLARED = True
    return L


def supported(group, hidden, top_k, dtype):
    return (dtype == torch.bfloat16 and hidden % 8 == 0 and top_k in (1, 2, 4, 6, 8) and _symm.is_supported(group)
            and hasattr(K.lib(), "dsb200_moe_dispatch"))


@dataclass
class DispatchPlan:
    """Everything both directions need for one MoE layer call (built from the gathered ``[ep, E]`` count matrix)."""
    T: int
    k: int
    order: torch.Tensor        # [T*k] int64: slot (t*k+j) at each expert-sorted position
    sorted_tok: torch.Tensor   # [T*k] int32: token row at each expert-sorted position
    segs: torch.Tensor         # [nseg, 4] int32 device (src_rank, src_start, dst_start, n) for MY experts
    nseg: int
    n_rows: int                # rows I receive
    local_counts: List[int]    # rows per local expert
    slot_rank: torch.Tensor    # [T*k] int32: expert rank of slot (t, j)
    slot_row: torch.Tensor     # [T*k] int32: row of that slot in the expert rank's buffers
    cap_tokens: int            # max over ranks of their token count
    cap_rows: int              # max over ranks of their received rows


def build_plan(top_i, num_experts, ep_size, ep_rank, ep_group):
    """``top_i`` [T, k] int64 expert choices of this rank's tokens."""
    T, k = top_i.shape
    dev = top_i.device
    flat_e = top_i.reshape(-1)
    order = torch.argsort(flat_e, stable=True)
    sorted_tok = (order // k).to(torch.int32)
    counts = torch.bincount(flat_e, minlength=num_experts)
    C = torch.empty(ep_size, num_experts, dtype=counts.dtype, device=dev)
    dist.all_gather_into_tensor(C.view(-1), counts, group=ep_group)
    Ch = C.tolist()  # the one host sync of the layer
    E_local = num_experts // ep_size
    # ---- my receive side: segments ordered (local expert, source rank) ----
    segs, dst, local_counts = [], 0, []
    for el in range(E_local):
        e = ep_rank * E_local + el
        tot = 0
        for p in range(ep_size):
            n = Ch[p][e]
            if n > 0:
                segs.append((p, sum(Ch[p][:e]), dst, n))
                dst += n
                tot += n
        local_counts.append(tot)
    n_rows = dst
    if not segs:
        segs = [(0, 0, 0, 0)]
    segs_t = torch.tensor(segs, dtype=torch.int32).to(dev, non_blocking=True)
    # ---- my send side: where does slot (t, j) sit in its expert rank's buffer? ----
    S_mine = torch.cumsum(C[ep_rank], 0) - C[ep_rank]                                   # exclusive prefix over experts
    before_me = C[:ep_rank].sum(0) if ep_rank > 0 else torch.zeros_like(C[0])          # rows from lower ranks, per e
    tot_e = C.sum(0).view(ep_size, E_local)
    base = (torch.cumsum(tot_e, 1) - tot_e).reshape(-1)                                 # first row of e on its rank
    inv = torch.empty_like(order)
    inv[order] = torch.arange(order.numel(), device=dev)
    slot_row = (base[flat_e] + before_me[flat_e] + (inv - S_mine[flat_e])).to(torch.int32)
    slot_rank = (flat_e // E_local).to(torch.int32)
    recv_per_rank = [sum(Ch[p][r * E_local + el] for p in range(ep_size) for el in range(E_local)) for r in range(ep_size)]
    return DispatchPlan(T=T, k=k, order=order, sorted_tok=sorted_tok, segs=segs_t, nseg=len(segs), n_rows=n_rows,
                        local_counts=local_counts, slot_rank=slot_rank.contiguous(), slot_row=slot_row.contiguous(),
                        cap_tokens=max(sum(row) for row in Ch) // k, cap_rows=max(recv_per_rank))


class FusedMoEComm:
    """Symmetric staging buffers + epochs of one expert-parallel group."""

    def __init__(self, group, hidden):
        self.ctx = _symm.SymmContext.get(group)
        self.group, self.hidden = group, hidden
        self.rank, self.world = self.ctx.rank, self.ctx.world
        self.tok_buf = self.row_buf = self.idx_buf = self.scale_buf = None
        self.cap_tokens = self.cap_rows = self.cap_slots = 0
        self.last_d, self.last_c = 0, 0
        self.tickets = torch.zeros(4, dtype=torch.int32, device=self.ctx.device)
        self._null = (ctypes.c_uint64 * 16)(*([0] * 16))

    @staticmethod
    def get(group, hidden):
        key = (id(group), hidden)
        if key not in _COMMS:
            _COMMS[key] = FusedMoEComm(group, hidden)
        return _COMMS[key]

    def _ensure(self, plan: DispatchPlan):
        """Grow the symmetric buffers (collective: every rank derives the same sizes from the shared count matrix)."""
        slots = plan.cap_tokens * plan.k
        if plan.cap_tokens > self.cap_tokens or plan.cap_rows > self.cap_rows or slots > self.cap_slots:
            torch.cuda.synchronize()
            dist.barrier(group=self.group)  # nobody may still be reading the old buffers
            self.cap_tokens = max(self.cap_tokens, int(plan.cap_tokens * 1.25) + 8)
            self.cap_rows = max(self.cap_rows, int(plan.cap_rows * 1.25) + 8)
            self.cap_slots = max(self.cap_slots, int(slots * 1.25) + 8)
            esz = 2
            self.tok_buf = self.ctx.alloc(self.cap_tokens * self.hidden * esz, want_multicast=False)
            self.row_buf = self.ctx.alloc(self.cap_rows * self.hidden * esz, want_multicast=False)
            self.idx_buf = self.ctx.alloc(self.cap_slots * 4, want_multicast=False)
            self.scale_buf = self.ctx.alloc(self.cap_slots * 4, want_multicast=False)
            self.last_d = self.last_c = 0

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    # ---- kernels -----------------------------------------------------------------------------------
    def dispatch(self, src, idx, scale, plan: DispatchPlan):
        """src [T, H] bf16 token rows of this rank; idx int32 [T*k]; scale fp32 [T*k] or None -> [plan.n_rows, H]."""
        self._ensure(plan)
        c = self.ctx
        c._wait(CH_D_DONE, self.last_d)  # peers finished reading my staging buffers of the previous dispatch
        n_tok = src.shape[0]
        self.tok_buf.view(n_tok * self.hidden, torch.bfloat16).view(n_tok, self.hidden).copy_(src)
        self.idx_buf.view(idx.numel(), torch.int32).copy_(idx)
        if scale is not None:
            self.scale_buf.view(scale.numel(), torch.float32).copy_(scale)
        out = torch.empty((plan.n_rows, self.hidden), dtype=torch.bfloat16, device=src.device)
        ep = c._next(CH_D_READY)
        rc = _lib().dsb200_moe_dispatch(self.tok_buf.ptrs, self.idx_buf.ptrs,
                                        self.scale_buf.ptrs if scale is not None else self._null, c.flags.ptrs, self.rank,
                                        self.world, K._ptr(plan.segs), plan.nseg, plan.n_rows, self.hidden, K._ptr(out),
                                        ctypes.c_void_p(self.tickets.data_ptr()), CH_D_READY, CH_D_DONE, ep, 0,
                                        self._stream())
        if rc < 0:
            raise RuntimeError(f"moe_dispatch failed: {K.last_error()}")
        K._LAUNCHES += 1
        self.last_d = ep
        return out

    def combine(self, rows, weights, plan: DispatchPlan, keep_rows=False):
        """rows [n_rows, H] bf16 expert-side rows of this rank; weights fp32 [T*k] or None -> ([T, H], kept rows)."""
        self._ensure(plan)
        c = self.ctx
        c._wait(CH_C_DONE, self.last_c)
        n = rows.shape[0]
        if n > 0:
            self.row_buf.view(n * self.hidden, torch.bfloat16).view(n, self.hidden).copy_(rows)
        out = torch.empty((plan.T, self.hidden), dtype=torch.bfloat16, device=rows.device)
        kept = torch.empty((plan.T * plan.k, self.hidden), dtype=torch.bfloat16, device=rows.device) if keep_rows else None
        ep = c._next(CH_C_READY)
        rc = _lib().dsb200_moe_combine(self.row_buf.ptrs, c.flags.ptrs, self.rank, self.world, K._ptr(plan.slot_rank),
                                       K._ptr(plan.slot_row), K._ptr(weights), plan.T, plan.k, self.hidden, K._ptr(out),
                                       K._ptr(kept), ctypes.c_void_p(self.tickets.data_ptr() + 4), CH_C_READY, CH_C_DONE,
                                       ep, 0, self._stream())
        if rc < 0:
            raise RuntimeError(f"moe_combine failed: {K.last_error()}")
        K._LAUNCHES += 1
        self.last_c = ep
        return out, kept


class _Dispatch(torch.autograd.Function):

    @staticmethod
    def forward(ctx, x, comm, plan):
        ctx.comm, ctx.plan = comm, plan
        return comm.dispatch(x.contiguous(), plan.sorted_tok, None, plan)

    @staticmethod
    def backward(ctx, g):
        dx, _ = ctx.comm.combine(g.contiguous(), None, ctx.plan)
        return dx, None, None


class _Combine(torch.autograd.Function):

    @staticmethod
    def forward(ctx, y, top_w, comm, plan):
        w = top_w.reshape(-1).to(torch.float32).contiguous()
        out, kept = comm.combine(y.contiguous(), w, plan, keep_rows=top_w.requires_grad or y.requires_grad)
        ctx.comm, ctx.plan = comm, plan
        ctx.save_for_backward(kept if kept is not None else torch.empty(0), w)
        ctx.w_shape, ctx.w_dtype = top_w.shape, top_w.dtype
        return out

    @staticmethod
    def backward(ctx, g):
        kept, w = ctx.saved_tensors
        plan = ctx.plan
        g = g.contiguous()
        d_w = None
        if ctx.needs_input_grad[1]:
            d_w = torch.einsum("tkh,th->tk", kept.view(plan.T, plan.k, -1).float(), g.float())
            d_w = d_w.reshape(ctx.w_shape).to(ctx.w_dtype)
        d_y = None
        if ctx.needs_input_grad[0]:
            d_y = ctx.comm.dispatch(g, plan.sorted_tok, w.index_select(0, plan.order).contiguous(), plan)
        return d_y, d_w, None, None


def fused_dispatch(x, comm, plan):
    """x [T, H] -> expert-contiguous rows [plan.n_rows, H] of the tokens routed to this rank's experts."""
    return _Dispatch.apply(x, comm, plan)


def fused_combine(expert_out, top_w, comm, plan):
    """expert_out [plan.n_rows, H], top_w [T, k] -> [T, H] (sum over the k experts of each token, weighted)."""
    return _Combine.apply(expert_out, top_w, comm, plan)
