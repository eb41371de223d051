"""ctypes binding to the in-tree sm_100a kernel library (``deepspeed_b200/ops/lib/libdsb200.so``).

The library exposes a plain C ABI (declared in ``include/dsb200.h``): every launcher takes raw device pointers,
sizes and a ``cudaStream_t``. Python passes ``tensor.data_ptr()`` and the current torch stream, so the kernels
run in-order with the surrounding PyTorch work and no torch C++ headers are needed to build (the whole library
cross-compiles in seconds with ``nvcc -gencode arch=compute_100a,code=sm_100a``).

Replaces the reference's per-op JIT builders (``op_builder/builder.py`` ``OpBuilder.load()``, 5.9 k LoC): there
is one target architecture, so one shared object.

On a machine with a GPU the library is REQUIRED: a missing/unbuildable library raises instead of silently
falling back to eager PyTorch.
"""
import ctypes
import os
import threading

import torch

_LIB = None
_LIB_ERR = None
_LOCK = threading.Lock()
LIB_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib")
LIB_PATH = os.path.join(LIB_DIR, "libdsb200.so")

c_void_p, c_int, c_int64, c_float, c_double = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_double


def _load():
    global _LIB, _LIB_ERR
    if _LIB is not None or _LIB_ERR is not None:
        return _LIB
    with _LOCK:
        if _LIB is not None:
            return _LIB
        try:
            if not os.path.exists(LIB_PATH):
                from .builder import build_cuda_library
                build_cuda_library(verbose=False)
            _LIB = ctypes.CDLL(LIB_PATH)
            _declare(_LIB)
        except Exception as e:  # noqa: BLE001
            _LIB_ERR = e
            _LIB = None
    return _LIB


def available():
    """True when CUDA is usable and the kernel library is loaded. Raises on a GPU box without the library."""
    if not torch.cuda.is_available():
        return False
    lib = _load()
    if lib is None:
        raise RuntimeError(f"deepspeed_b200: CUDA device present but kernel library failed to load/build: {_LIB_ERR}")
    return True


def lib():
    if not available():
        raise RuntimeError("deepspeed_b200 kernel library requires a CUDA device")
    return _LIB


def _stream():
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return c_void_p(t.data_ptr()) if t is not None else c_void_p(0)


_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def _dt(t):
    return _DT[t.dtype]


_LAUNCHES = 0
_KERNELS_PER_CALL = {"rmsnorm_bwd": 2, "layernorm_bwd": 3, "lamb": 2}


def launch_count():
    """Number of kernels of this library launched so far in this process."""
    return _LAUNCHES


def _check(rc, what):
    global _LAUNCHES
    if rc != 0:
        raise RuntimeError(f"dsb200 kernel '{what}' failed with code {rc}: {last_error()}")
    _LAUNCHES += _KERNELS_PER_CALL.get(what, 1)


def last_error():
    try:
        _LIB.dsb200_last_error.restype = ctypes.c_char_p
        return _LIB.dsb200_last_error().decode()
    except Exception:  # noqa: BLE001
        return "?"


def _declare(L):
    """Argument/return types for every entry point (mirrors include/dsb200.h)."""
    def sig(name, *argtypes):
        fn = getattr(L, name, None)
        if fn is not None:
            fn.argtypes = list(argtypes)
            fn.restype = c_int
    P, I, L64, F, D = c_void_p, c_int, c_int64, c_float, c_double
    sig("dsb200_sumsq", P, L64, I, P, P)
    sig("dsb200_multi_sumsq", P, I, P, P)
    sig("dsb200_multi_tensor_optim", I, P, I, P, I, F, P)
    sig("dsb200_lamb_step", P, P, P, P, P, L64, I, P, P, F, F, F, F, F, F, F, I, F, F, P)
    sig("dsb200_rmsnorm_fwd", P, P, P, P, P, P, I, I, F, I, P)
    sig("dsb200_rmsnorm_bwd", P, P, P, P, P, P, P, P, I, I, I, I, P)
    sig("dsb200_layernorm_fwd", P, P, P, P, P, P, P, I, I, F, I, P)
    sig("dsb200_layernorm_bwd", P, P, P, P, P, P, P, P, P, I, I, I, I, P)
    sig("dsb200_rope", P, P, P, P, I, I, I, L64, I, I, I, P)
    sig("dsb200_swiglu_fwd", P, P, P, L64, I, L64, I, P)
    sig("dsb200_swiglu_bwd", P, P, P, P, P, L64, I, L64, I, P)
    sig("dsb200_ce_fwd", P, P, P, P, I, I, L64, I, I, F, I, P)
    sig("dsb200_ce_fwd_ex", P, P, P, P, I, I, L64, I, I, F, P, I, P)
    sig("dsb200_scale_by_dev", P, L64, I, P, P)
    sig("dsb200_wo_gemv", P, P, P, P, P, P, I, I, I, I, I, L64, L64, P)
    sig("dsb200_wo_gemm_tc", P, P, P, P, P, P, P, I, I, I, I, I, L64, L64, P)
    sig("dsb200_ragged_kv_append_rope", P, P, P, P, P, P, P, P, P, I, L64, L64, L64, I, I, I, I, I, I, I, P)
    sig("dsb200_ragged_paged_attention", P, P, P, P, P, P, P, P, P, I, L64, I, I, I, I, I, I, F, P)
    sig("dsb200_move_rows", P, P, P, I, L64, L64, L64, L64, L64, I, P)
    sig("dsb200_scatter_add_rows", P, P, P, I, P, L64, I, L64, L64, L64, I, I, P)
    sig("dsb200_sort_rows_i32", P, L64, I, P)


# ---------------------------------------------------------------------------------------------------------
# reductions
# ---------------------------------------------------------------------------------------------------------
_SCRATCH = {}
_SUMSQ_SCRATCH = {}


def _scratch(device, n_float):
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    buf = _SCRATCH.get(key)
    if buf is None or buf.numel() < n_float:
        buf = torch.zeros(max(n_float, 8192), dtype=torch.float32, device=device)
        _SCRATCH[key] = buf
    return buf


def sumsq(t):
    """fp32 sum of squares of one contiguous tensor (two-stage deterministic reduction)."""
    t = t.contiguous()
    out = torch.empty((), dtype=torch.float32, device=t.device)
    # dedicated buffer: the ticket word (index 1024) must stay zero between launches, so this scratch must never be
    # shared with kernels that scribble partial sums over it (norm backward, LAMB use _scratch)
    key = (t.device.index, torch.cuda.current_stream().cuda_stream)
    scratch = _SUMSQ_SCRATCH.get(key)
    if scratch is None:
        scratch = _SUMSQ_SCRATCH[key] = torch.zeros(2048, dtype=torch.float32, device=t.device)
    _check(lib().dsb200_sumsq(_ptr(t), t.numel(), _dt(t), _ptr(scratch), _ptr(out), _stream()), "sumsq")
    return out


def multi_tensor_sumsq(tensors):
    total = None
    for t in tensors:
        s = sumsq(t)
        total = s if total is None else total + s
    return total


# ---------------------------------------------------------------------------------------------------------
# multi-tensor optimizers
# ---------------------------------------------------------------------------------------------------------
CHUNK = 65536
_KIND = {"adam": 0, "lion": 1, "adagrad": 2, "sgd": 3}


class ChunkTable:
    """Device-resident table of work chunks: (p, g, m, v, lp, n, group, gdt|ldt|sdt)."""

    def __init__(self, dev_table, n_chunks, keepalive):
        self.dev_table, self.n_chunks, self.keepalive = dev_table, n_chunks, keepalive


def build_chunk_table(items, states):
    import numpy as np
    rows = []
    for (gi, p, g, lp), sts in zip(items, states):
        assert p.dtype == torch.float32 or p.dtype in _DT
        n = p.numel()
        m, v = sts[0], (sts[1] if len(sts) > 1 else None)  # [first-moment slot, second-moment slot]; either may be None
        esz = {0: 4, 1: 2, 2: 2}
        pdt, gdt = _dt(p), _dt(g)
        sdt = _dt(m if m is not None else v)
        ldt = _dt(lp) if lp is not None else 0
        for off in range(0, n, CHUNK):
            cnt = min(CHUNK, n - off)
            rows.append((p.data_ptr() + off * esz[pdt], g.data_ptr() + off * esz[gdt],
                         (m.data_ptr() + off * esz[sdt]) if m is not None else 0,
                         (v.data_ptr() + off * esz[sdt]) if v is not None else 0,
                         (lp.data_ptr() + off * esz[ldt]) if lp is not None else 0, cnt, gi,
                         pdt | (gdt << 4) | (sdt << 8) | (ldt << 12)))
    arr = np.zeros(len(rows), dtype=np.dtype([("p", "u8"), ("g", "u8"), ("m", "u8"), ("v", "u8"), ("lp", "u8"),
                                              ("n", "i4"), ("group", "i4"), ("dt", "i4"), ("pad", "i4")]))
    for i, r in enumerate(rows):
        arr[i] = r + (0, )
    host = torch.from_numpy(arr.view(np.uint8).copy())
    dev = host.to(items[0][1].device)
    return ChunkTable(dev, len(rows), [items, states])


def multi_tensor_optim(kind, table, hyper, grad_scale):
    import numpy as np
    assert len(hyper) <= 64, "at most 64 param groups per launch"
    h = np.zeros((len(hyper), 8), dtype=np.float32)
    for i, d in enumerate(hyper):
        h[i] = (d["lr"], d["beta1"], d["beta2"], d["eps"], d["wd"], d["bc1"], d["bc2"], float(d["mode"]))
    _check(
        lib().dsb200_multi_tensor_optim(_KIND[kind], _ptr(table.dev_table), table.n_chunks,
                                        h.ctypes.data_as(c_void_p), len(hyper), float(grad_scale), _stream()),
        f"multi_tensor_{kind}")


def lamb_step(p, g, m, v, lp, h, coeffs_out):
    scratch = _scratch(p.device, 4096 + 8)
    _check(
        lib().dsb200_lamb_step(_ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(lp), p.numel(), (_dt(lp) if lp is not None else 0),
                               _ptr(scratch), _ptr(scratch[4096:]), h["lr"], h["beta1"], h["beta2"], h["eps"], h["wd"],
                               h["bc1"], h["bc2"], int(h["mode"]), h["max_coeff"], h["min_coeff"], _stream()), "lamb")
    coeffs_out.append(scratch[4096 + 2].clone())  # the scratch word is reused by the next launch


# ---------------------------------------------------------------------------------------------------------
# norms
# ---------------------------------------------------------------------------------------------------------
def _norm_grid(rows, cols):
    return min(rows, 148 * 4)


def rmsnorm_fwd(x2, res2, weight, eps):
    rows, cols = x2.shape
    y = torch.empty_like(x2)
    res_out = torch.empty_like(x2) if res2 is not None else None
    rstd = torch.empty(rows, dtype=torch.float32, device=x2.device)
    _check(lib().dsb200_rmsnorm_fwd(_ptr(x2), _ptr(res2), _ptr(weight), _ptr(y), _ptr(res_out), _ptr(rstd), rows, cols,
                                    float(eps), _dt(x2), _stream()), "rmsnorm_fwd")
    return y, res_out, rstd


def rmsnorm_bwd(dy2, x2, weight, rstd, dres2):
    rows, cols = x2.shape
    dx = torch.empty_like(x2)
    dw = torch.empty_like(weight)
    scratch = _scratch(x2.device, _norm_grid(rows, cols) * cols)
    _check(lib().dsb200_rmsnorm_bwd(_ptr(dy2), _ptr(x2), _ptr(weight), _ptr(rstd), _ptr(dres2), _ptr(dx), _ptr(dw),
                                    _ptr(scratch), rows, cols, 0, _dt(x2), _stream()), "rmsnorm_bwd")
    return dx, dw


def layernorm_fwd(x2, weight, bias, eps):
    rows, cols = x2.shape
    y = torch.empty_like(x2)
    stats = torch.empty(rows * 2, dtype=torch.float32, device=x2.device)
    _check(lib().dsb200_layernorm_fwd(_ptr(x2), c_void_p(0), _ptr(weight), _ptr(bias), _ptr(y), c_void_p(0),
                                      _ptr(stats), rows, cols, float(eps), _dt(x2), _stream()), "layernorm_fwd")
    return y, stats


def layernorm_bwd(dy2, x2, weight, stats, has_bias):
    rows, cols = x2.shape
    dx = torch.empty_like(x2)
    dw = torch.empty_like(weight)
    db = torch.empty_like(weight) if has_bias else None
    scratch = _scratch(x2.device, 2 * _norm_grid(rows, cols) * cols)
    _check(lib().dsb200_layernorm_bwd(_ptr(dy2), _ptr(x2), _ptr(weight), _ptr(stats), c_void_p(0), _ptr(dx), _ptr(dw),
                                      _ptr(db), _ptr(scratch), rows, cols, 0, _dt(x2), _stream()), "layernorm_bwd")
    return dx, dw, db


# ---------------------------------------------------------------------------------------------------------
# rope / swiglu / cross-entropy
# ---------------------------------------------------------------------------------------------------------
def rope_(x, cos, sin, pos, seq_len, backward=False):
    """In-place half-rotation RoPE on contiguous x [B, S, H, D]."""
    B, S, H, D = x.shape
    assert x.is_contiguous() and cos.dtype == torch.float32 and cos.shape[-1] == D // 2
    _check(lib().dsb200_rope(_ptr(x), _ptr(cos), _ptr(sin), _ptr(pos), B * S, H, D, H * D, seq_len,
                             1 if backward else 0, _dt(x), _stream()), "rope")
    return x


def rope_strided_(x, cos, sin, pos, rows, n_heads, head_dim, row_stride, seq_len, backward=False):
    """In-place half-rotation RoPE on the first ``n_heads`` heads of every row of ``x`` (rows ``row_stride``
    elements apart): the q and k heads of a fused QKV projection in one launch."""
    assert x.is_contiguous() and cos.dtype == torch.float32 and cos.shape[-1] == head_dim // 2
    _check(lib().dsb200_rope(_ptr(x), _ptr(cos), _ptr(sin), _ptr(pos), rows, n_heads, head_dim, row_stride, seq_len,
                             1 if backward else 0, _dt(x), _stream()), "rope")
    return x


def swiglu_fwd(gate_up):
    inter = gate_up.shape[-1] // 2
    rows = gate_up.numel() // (2 * inter)
    y = torch.empty(gate_up.shape[:-1] + (inter, ), dtype=gate_up.dtype, device=gate_up.device)
    esz = gate_up.element_size()
    _check(lib().dsb200_swiglu_fwd(c_void_p(gate_up.data_ptr()), c_void_p(gate_up.data_ptr() + inter * esz), _ptr(y),
                                   rows, inter, 2 * inter, _dt(gate_up), _stream()), "swiglu_fwd")
    return y


def swiglu_bwd(dy, gate_up):
    inter = gate_up.shape[-1] // 2
    rows = gate_up.numel() // (2 * inter)
    dgu = torch.empty_like(gate_up)
    esz = gate_up.element_size()
    _check(lib().dsb200_swiglu_bwd(_ptr(dy), c_void_p(gate_up.data_ptr()), c_void_p(gate_up.data_ptr() + inter * esz),
                                   c_void_p(dgu.data_ptr()), c_void_p(dgu.data_ptr() + inter * esz), rows, inter,
                                   2 * inter, _dt(gate_up), _stream()), "swiglu_bwd")
    return dgu


def cross_entropy_inplace_grad(logits, target, ignore_index, grad_scale, grad_scale_dev=None):
    """Per-row loss (fp32); ``logits`` is overwritten with ``(softmax - onehot) * grad_scale`` (times the fp32 device
    scalar ``grad_scale_dev`` when given)."""
    rows, vocab = logits.shape
    assert logits.stride(1) == 1
    loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    _check(lib().dsb200_ce_fwd_ex(_ptr(logits), _ptr(target), _ptr(loss), c_void_p(0), rows, vocab, logits.stride(0),
                                  int(ignore_index), 1, float(grad_scale), _ptr(grad_scale_dev), _dt(logits), _stream()),
           "ce_fwd")
    return loss


def scale_by_device_scalar_(x, scale_dev):
    """``x *= scale_dev`` (0-dim fp32 device tensor) without a host sync; nothing is touched when the scalar is 1."""
    assert x.is_contiguous() and scale_dev.dtype == torch.float32
    _check(lib().dsb200_scale_by_dev(_ptr(x), x.numel(), _dt(x), _ptr(scale_dev), _stream()), "scale_by_dev")
    return x


def cross_entropy_fwd(logits, target, ignore_index):
    rows, vocab = logits.shape
    loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
    _check(lib().dsb200_ce_fwd(_ptr(logits), _ptr(target), _ptr(loss), _ptr(lse), rows, vocab, logits.stride(0),
                               int(ignore_index), 0, 1.0, _dt(logits), _stream()), "ce_fwd")
    return loss, lse


# ---------------------------------------------------------------------------------------------------------
# row gather / scatter by index (csrc/token_ops.cu)
# ---------------------------------------------------------------------------------------------------------
def _idx_bytes(idx):
    assert idx.dtype in (torch.int32, torch.int64) and idx.is_contiguous()
    return idx.element_size()


def gather_rows(src2, idx, out=None):
    """``out[i] = src2[idx[i]]`` for a 2-D ``src2`` with contiguous rows; out-of-range indices leave ``out[i]``
    untouched (pass a zeroed ``out`` for MoE capacity drops)."""
    assert src2.dim() == 2 and src2.stride(1) == 1
    n = idx.numel()
    if out is None:
        out = torch.empty((n, src2.shape[1]), dtype=src2.dtype, device=src2.device)
    esz = src2.element_size()
    _check(lib().dsb200_move_rows(_ptr(src2), _ptr(out), _ptr(idx), _idx_bytes(idx), n, src2.shape[1] * esz,
                                  src2.stride(0) * esz, out.stride(0) * esz, src2.shape[0], 0, _stream()),
           "move_rows")
    return out


def scatter_rows_(dst2, idx, src2):
    """``dst2[idx[i]] = src2[i]`` (indices unique); returns ``dst2``."""
    assert dst2.dim() == 2 and dst2.stride(1) == 1 and src2.stride(1) == 1 and dst2.dtype == src2.dtype
    esz = src2.element_size()
    _check(lib().dsb200_move_rows(_ptr(src2), _ptr(dst2), _ptr(idx), _idx_bytes(idx), idx.numel(),
                                  src2.shape[1] * esz, src2.stride(0) * esz, dst2.stride(0) * esz, dst2.shape[0], 1,
                                  _stream()), "move_rows")
    return dst2


def scatter_add_rows_(dst2, idx, src2, weights=None, unique=False):
    """``dst2[idx[i]] += weights[i] * src2[i]``; atomics unless the indices are known to be unique."""
    assert dst2.dtype == src2.dtype and dst2.stride(1) == 1 and src2.stride(1) == 1
    w = _ptr(weights) if weights is not None else c_void_p(0)
    if weights is not None:
        assert weights.dtype == torch.float32 and weights.is_contiguous()
    _check(lib().dsb200_scatter_add_rows(_ptr(src2), _ptr(dst2), _ptr(idx), _idx_bytes(idx), w, idx.numel(),
                                         src2.shape[1], src2.stride(0), dst2.stride(0), dst2.shape[0],
                                         0 if unique else 1, _dt(src2), _stream()), "scatter_add_rows")
    return dst2


def sort_rows_i32_(idx2):
    """Ascending in-place sort of every row of an int32 [rows, n<=8192] tensor."""
    assert idx2.dtype == torch.int32 and idx2.is_contiguous()
    n = idx2.shape[-1]
    _check(lib().dsb200_sort_rows_i32(_ptr(idx2), idx2.numel() // n, n, _stream()), "sort_rows_i32")
    return idx2


# ---------------------------------------------------------------------------------------------------------
# tcgen05 GEMM
# ---------------------------------------------------------------------------------------------------------
def gemm_bf16_nt(a, b, out=None, max_ctas=0, block_n=0):
    """``out[M,N] = a[M,K] @ b[N,K].T`` (bf16 in/out, fp32 accumulate in TMEM) on the hand-written tcgen05 kernel."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.dim() == 2 and b.dim() == 2
    assert a.stride(1) == 1 and b.stride(1) == 1 and a.shape[1] == b.shape[1]
    M, Kd = a.shape
    N = b.shape[0]
    if out is None:
        out = torch.empty(M, N, dtype=torch.bfloat16, device=a.device)
    L = lib()
    if not hasattr(L, "_gemm_declared"):
        L.dsb200_gemm_bf16_nt.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_int64, c_int64,
                                          c_int, c_int, c_void_p]
        L.dsb200_gemm_bf16_nt.restype = c_int
        L._gemm_declared = True
    _check(L.dsb200_gemm_bf16_nt(_ptr(a), _ptr(b), _ptr(out), M, N, Kd, a.stride(0), b.stride(0), out.stride(0),
                                 int(max_ctas), int(block_n), _stream()), "gemm_bf16_nt")
    return out


# ---------------------------------------------------------------------------------------------------------
# FastGen ragged ops (csrc/ragged_ops.cu)
# ---------------------------------------------------------------------------------------------------------
def ragged_kv_append_rope(q, k, v, kv_layer, cos, sin, tok_seq, tok_pos, block_table, rotary_dim):
    """RoPE on q [T, Hq, D] in place and on k, then k / v [T, Hkv, D] are appended to the blocked cache of one layer
    ``kv_layer`` [num_blocks, 2, block_size, Hkv, D]. ``rotary_dim`` 0: positions are not rotary."""
    T, Hq, D = q.shape
    Hkv = k.shape[1]
    assert q.stride(2) == 1 and q.stride(1) == D and k.stride(1) == D and v.stride(1) == D and kv_layer.is_contiguous()
    _check(lib().dsb200_ragged_kv_append_rope(
        _ptr(q), _ptr(k), _ptr(v), _ptr(kv_layer), _ptr(cos), _ptr(sin), _ptr(tok_seq), _ptr(tok_pos), _ptr(block_table),
        T, q.stride(0), k.stride(0), v.stride(0), Hq, Hkv, D, int(rotary_dim), kv_layer.shape[2], block_table.shape[1],
        _dt(q), _stream()), "ragged_kv_append_rope")


_PA_SCRATCH = {}


def ragged_paged_attention(q, kv_layer, tok_seq, tok_pos, block_table, scale, max_context, out=None, tok_row=None):
    """Every listed token attends to keys [0, pos] of its sequence in the blocked cache.
    q [T, Hq, D] (bf16, D in {64, 128}) -> out [T, Hq, D]. ``tok_row`` (int32) restricts the launch to a subset of the
    rows of q / out (``tok_seq`` / ``tok_pos`` then describe that subset). The key range is split over several CTAs
    (flash-decoding) when the batch alone cannot fill the GPU."""
    Hq, D = q.shape[1], q.shape[2]
    T = q.shape[0] if tok_row is None else tok_row.numel()
    Hkv = kv_layer.shape[3]
    if out is None:
        out = torch.empty((q.shape[0], Hq, D), dtype=q.dtype, device=q.device)
    tiles = max(1, (int(max_context) + 31) // 32)
    splits = 1
    if T * Hkv < 2 * 148:
        splits = max(1, min(tiles // 4, (4 * 148) // max(1, T * Hkv), 32))
    pa = pm = None
    if splits > 1:
        key = (q.device, T * Hq * splits * (D + 2))
        buf = _PA_SCRATCH.get(key)
        if buf is None:
            _PA_SCRATCH.clear()
            buf = _PA_SCRATCH[key] = torch.empty(T * Hq * splits * (D + 2), dtype=torch.float32, device=q.device)
        pa, pm = buf[:T * Hq * splits * D], buf[T * Hq * splits * D:]
    rc = lib().dsb200_ragged_paged_attention(_ptr(q), _ptr(kv_layer), _ptr(out), _ptr(tok_seq), _ptr(tok_pos),
                                             _ptr(block_table), _ptr(tok_row), _ptr(pa), _ptr(pm), T, q.stride(0), Hq,
                                             Hkv, D,
                                             kv_layer.shape[2], block_table.shape[1], splits, float(scale), _stream())
    if rc < 0:
        raise RuntimeError(f"ragged_paged_attention failed: {last_error()}")
    global _LAUNCHES
    _LAUNCHES += 1
    return out


def ragged_attention_supported(q, head_dim, n_heads, n_kv_heads):
    return (available() and q.is_cuda and q.dtype == torch.bfloat16 and head_dim in (64, 128)
            and n_heads % n_kv_heads == 0 and n_heads // n_kv_heads in (1, 2, 4, 8))


# ---------------------------------------------------------------------------------------------------------
# weight-only quantized decode GEMM (csrc/wo_gemm.cu)
# ---------------------------------------------------------------------------------------------------------
_WO_SCRATCH = {}


def wo_gemv_supported(x2, payload, group, q_bits, K_dim):
    return (available() and x2.is_cuda and x2.dtype == torch.bfloat16 and x2.shape[0] <= 8 and q_bits in (4, 8)
            and payload.dtype == torch.uint8 and K_dim % 256 == 0 and group % 16 == 0 and K_dim % group == 0
            and x2.stride(1) == 1 and x2.stride(0) % 8 == 0)


def wo_gemm_supported(x2, payload, group, q_bits, K_dim):
    """Shapes the tcgen05 weight-only GEMM serves: up to 64 tokens."""
    return (available() and x2.is_cuda and x2.dtype == torch.bfloat16 and 0 < x2.shape[0] <= 64 and q_bits in (4, 6, 8)
            and payload.dtype == torch.uint8 and K_dim % 64 == 0 and group % 64 == 0 and K_dim % group == 0
            and x2.stride(1) == 1 and x2.stride(0) % 8 == 0 and x2.data_ptr() % 16 == 0 and payload.data_ptr() % 16 == 0
            and (K_dim * q_bits // 8) % 16 == 0)


def wo_gemm(x2, payload, scales, N, K_dim, group, q_bits, bias=None):
    """``x2 [T <= 64, K] @ dequant(W [N, K])^T`` on the tensor cores (csrc/wo_gemm.cu, tcgen05 version): the quantized
    weights are streamed once by TMA, dequantised into the swizzled operand layout in shared memory and multiplied with
    the tokens as the N extent of the MMA. Returns bf16 [T, N]."""
    T = x2.shape[0]
    y = torch.empty((T, N), dtype=torch.bfloat16, device=x2.device)
    need = T * N + (T * K_dim + 1) // 2 + 64
    buf = _WO_SCRATCH.get(x2.device)
    if buf is None or buf.numel() < need:
        buf = _WO_SCRATCH[x2.device] = torch.empty(max(need, 64 * 32768), dtype=torch.float32, device=x2.device)
    y32, x16 = buf[:T * N], buf[(T * N + 3) // 4 * 4:]
    _check(lib().dsb200_wo_gemm_tc(_ptr(x2), _ptr(payload), _ptr(scales), _ptr(bias), _ptr(y), _ptr(y32), _ptr(x16), T, N,
                                   K_dim, int(group), int(q_bits), x2.stride(0), y.stride(0), _stream()), "wo_gemm_tc")
    return y


def wo_gemv(x2, payload, scales, N, K_dim, group, q_bits, bias=None):
    """``x2 [M <= 8, K] @ dequant(W [N, K])^T`` with W as FP8 (e4m3 bytes) / FP4 (e2m1 nibble pairs) + fp32 group
    scales; the weights are read once in their quantized form. Returns bf16 [M, N]."""
    M = x2.shape[0]
    y = torch.empty((M, N), dtype=torch.bfloat16, device=x2.device)
    buf = _WO_SCRATCH.get(x2.device)
    if buf is None or buf.numel() < M * N:
        buf = _WO_SCRATCH[x2.device] = torch.empty(max(M * N, 8 * 32768), dtype=torch.float32, device=x2.device)
    _check(lib().dsb200_wo_gemv(_ptr(x2), _ptr(payload), _ptr(scales), _ptr(bias), _ptr(y), _ptr(buf), M, N, K_dim,
                                int(group), int(q_bits), x2.stride(0), y.stride(0), _stream()), "wo_gemv")
    return y
