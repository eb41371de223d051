"""ZeRO++ quantization ops: qwZ (quantized weight all-gather) and qgZ (quantized gradient reduce).

Reference API: ``QuantizerBuilder().load()`` -> ``quantize / dequantize / swizzle_quant / quantized_reduction /
loco_*`` (``csrc/quantization/pt_binding.cpp:372-401``), ``CUDAQuantizer`` (``partition_parameters.py:846-890``),
``all_to_all_quant_reduce`` (``runtime/comm/coalesced_collectives.py:31-153``).
Native wire format here: block-scaled FP8 e4m3 (``cvt.rn.satfinite.e4m3x2``) with one fp32 scale per block; the int8 /
int4 symmetric formats are kept for parity. CUDA kernels: ``csrc/quant.cu``; CPU fallback in PyTorch (tests).
"""
import ctypes

import torch

from .. import comm as dist
from . import kernels as K

FP8_BLOCK = 512
_declared = False


def _lib():
    global _declared
    L = K.lib()
    if not _declared:
        P, I, L64, F, U = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float, ctypes.c_uint
        L.dsb200_fp8_block_quant.argtypes = [P, I, P, P, L64, I, I, P]
        L.dsb200_fp8_block_dequant.argtypes = [P, P, P, I, L64, I, I, P]
        L.dsb200_int_group_quant.argtypes = [P, I, P, P, L64, I, I, I, I, P]
        L.dsb200_int_group_dequant.argtypes = [P, P, P, I, L64, I, I, P]
        L.dsb200_dequant_reduce.argtypes = [P, P, P, I, L64, I, I, I, F, P]
        L.dsb200_minifloat_quant.argtypes = [P, P, P, I, L64, I, I, I, F, I, U, P]
        if hasattr(L, "dsb200_loco_group_quant"):
            L.dsb200_loco_group_quant.argtypes = [P, I, P, P, P, L64, I, I, F, I, I, P]
            L.dsb200_fake_quant.argtypes = [P, I, L64, I, I, I, I, U, P]
        _declared = True
    return L


def _cuda(x):
    return x.is_cuda and K.available()


# ---- FP8 block-scaled -------------------------------------------------------------------------------------
def fp8_block_quantize(x, block=FP8_BLOCK, fmt="e4m3"):
    """x (bf16/fp16, numel % block == 0) -> (uint8 payload, fp32 scales[numel/block])."""
    x = x.contiguous().view(-1)
    n = x.numel()
    assert n % block == 0 and block % 256 == 0
    f = 0 if fmt == "e4m3" else 1
    if _cuda(x):
        q = torch.empty(n, dtype=torch.uint8, device=x.device)
        s = torch.empty(n // block, dtype=torch.float32, device=x.device)
        K._check(_lib().dsb200_fp8_block_quant(K._ptr(x), K._dt(x), K._ptr(q), K._ptr(s), n, block, f, K._stream()),
                 "fp8_block_quant")
        return q, s
    fmax = 448.0 if f == 0 else 57344.0
    dt = torch.float8_e4m3fn if f == 0 else torch.float8_e5m2
    xb = x.float().view(-1, block)
    s = xb.abs().amax(dim=1) / fmax
    s = torch.where(s > 0, s, torch.ones_like(s))
    q = (xb / s[:, None]).clamp(-fmax, fmax).to(dt).view(torch.uint8).view(-1)
    return q, s


def fp8_block_dequantize(q, s, dtype=torch.bfloat16, block=FP8_BLOCK, fmt="e4m3", out=None):
    n = q.numel()
    f = 0 if fmt == "e4m3" else 1
    if out is None:
        out = torch.empty(n, dtype=dtype, device=q.device)
    if _cuda(q):
        K._check(_lib().dsb200_fp8_block_dequant(K._ptr(q), K._ptr(s), K._ptr(out), K._dt(out), n, block, f,
                                                 K._stream()), "fp8_block_dequant")
        return out
    dt = torch.float8_e4m3fn if f == 0 else torch.float8_e5m2
    out.copy_((q.view(dt).float().view(-1, block) * s[:, None]).view(-1).to(dtype))
    return out


# ---- integer group quantization (reference parity) ---------------------------------------------------------
def quantize(x, groups, bits=8, sym=True, swizzle=None):
    """Symmetric int8/int4 group quantization. Returns (int8 tensor [numel*bits/8], fp32 scales [groups])."""
    assert sym, "asymmetric quantization is not used by ZeRO++ and is not implemented"
    x = x.contiguous().view(-1)
    n = x.numel()
    assert n % groups == 0
    group = n // groups
    nodes, devs = swizzle if swizzle else (0, 0)
    if _cuda(x) and group % 256 == 0:
        q = torch.empty(n * bits // 8, dtype=torch.int8, device=x.device)
        s = torch.empty(groups, dtype=torch.float32, device=x.device)
        K._check(_lib().dsb200_int_group_quant(K._ptr(x), K._dt(x), K._ptr(q), K._ptr(s), n, group, bits, nodes, devs,
                                               K._stream()), "int_group_quant")
        return q, s
    qmax = 127.0 if bits == 8 else 7.0
    xg = x.float().view(groups, group)
    s = xg.abs().amax(dim=1) / qmax
    s = torch.where(s > 0, s, torch.ones_like(s))
    qi = torch.round(xg / s[:, None]).clamp(-qmax, qmax).to(torch.int8)
    if nodes * devs > 0:
        gpp = groups // (nodes * devs)
        qi = qi.view(nodes, devs, gpp, group).permute(1, 0, 2, 3).contiguous().view(groups, group)
        s = s.view(nodes, devs, gpp).permute(1, 0, 2).contiguous().view(-1)
    if bits == 4:
        flat = qi.view(-1).to(torch.int16) & 0xF
        qi = (flat[0::2] | (flat[1::2] << 4)).to(torch.uint8).view(torch.int8)
    return qi.view(-1), s


def dequantize(q, s, groups, bits=8, dtype=torch.bfloat16, out=None):
    n = q.numel() * 8 // bits
    group = n // groups
    if out is None:
        out = torch.empty(n, dtype=dtype, device=q.device)
    if _cuda(q) and n % 8 == 0:
        K._check(_lib().dsb200_int_group_dequant(K._ptr(q), K._ptr(s), K._ptr(out), K._dt(out), n, group, bits,
                                                 K._stream()), "int_group_dequant")
        return out
    if bits == 4:
        u = q.view(torch.uint8).to(torch.int16)
        lo, hi = u & 0xF, (u >> 4) & 0xF
        vals = torch.stack([lo, hi], dim=1).view(-1)
        vals = ((vals ^ 8) - 8).float()
    else:
        vals = q.float()
    out.copy_((vals.view(groups, group) * s[:, None]).view(-1).to(dtype))
    return out


def loco_quantize(x, err, groups, bits=8, beta=0.8, swizzle=None):
    """LoCo error-feedback group quantization (reference ``loco_swizzled_quant``): quantizes ``x + beta * err`` and
    replaces ``err`` (fp32, shape of ``x``, updated in place) by what the quantization lost. Returns (q, scales)."""
    x = x.contiguous().view(-1)
    n = x.numel()
    assert n % groups == 0 and err.numel() == n and err.dtype == torch.float32 and err.is_contiguous()
    group = n // groups
    nodes, devs = swizzle if swizzle else (0, 0)
    if _cuda(x) and group % 256 == 0 and hasattr(_lib(), "dsb200_loco_group_quant"):
        q = torch.empty(n * bits // 8, dtype=torch.int8, device=x.device)
        s = torch.empty(groups, dtype=torch.float32, device=x.device)
        K._check(_lib().dsb200_loco_group_quant(K._ptr(x), K._dt(x), K._ptr(err), K._ptr(q), K._ptr(s), n, group, bits,
                                                float(beta), nodes, devs, K._stream()), "loco_group_quant")
        return q, s
    v = x.float() + beta * err.view(-1)
    q, s = quantize(v, groups, bits, True, swizzle=None)
    err.view(-1).copy_(v - dequantize(q, s, groups, bits, dtype=torch.float32))
    if nodes * devs > 0:
        q, s = quantize(v, groups, bits, True, swizzle=swizzle)
    return q, s


def fake_quantize(x, groups, bits=8, symmetric=True, stochastic=False, seed=0):
    """MoQ quantize-dequantize of ``x`` in place (reference ``fake_quantizer.cu`` ``ds_quantize`` / ``ds_sr_quantize``
    and the ``_asym`` variants). Returns ``x``."""
    flat = x.view(-1)
    n = flat.numel()
    assert n % groups == 0
    group = n // groups
    if _cuda(x) and x.is_contiguous() and group % 256 == 0 and x.dtype in (torch.bfloat16, torch.float16) and \
            hasattr(_lib(), "dsb200_fake_quant"):
        K._check(_lib().dsb200_fake_quant(K._ptr(x), K._dt(x), n, group, int(bits), int(bool(symmetric)),
                                          int(bool(stochastic)), int(seed) & 0xFFFFFFFF, K._stream()), "fake_quant")
        return x
    v = flat.float().view(groups, group)
    levels = float(1 << bits)
    if symmetric:
        amax = v.abs().amax(dim=1, keepdim=True)
        scale = torch.where(amax > 0, 2 * amax / levels, torch.ones_like(amax))
        zero, lo, hi = torch.zeros_like(amax), -levels / 2, levels / 2 - 1
    else:
        mn, mx = v.amin(dim=1, keepdim=True), v.amax(dim=1, keepdim=True)
        scale = torch.where(mx > mn, (mx - mn) / (levels - 1), torch.ones_like(mx))
        zero, lo, hi = mn, 0.0, levels - 1
    t = (v - zero) / scale
    if stochastic:
        g = torch.Generator(device=t.device).manual_seed(int(seed))
        t = torch.floor(t + torch.rand(t.shape, generator=g, device=t.device))
    else:
        t = torch.round(t)
    flat.copy_((t.clamp(lo, hi) * scale + zero).view(-1).to(x.dtype))
    return x


def swizzle_quant(x, groups, bits, nodes, devices_per_node):
    """qgZ stage 1: quantize and re-order partitions ``[node, dev] -> [dev, node]``."""
    return quantize(x, groups, bits, True, swizzle=(nodes, devices_per_node))


def quantized_reduction(q, s, in_groups, out_groups, bits, nchunks, dtype=torch.bfloat16, requantize_bits=None,
                        average=True):
    """qgZ stage 2: dequantize ``nchunks`` received chunks, reduce in fp32, optionally re-quantize."""
    n_total = q.numel() * 8 // bits
    chunk = n_total // nchunks
    group = n_total // in_groups
    post = 1.0 / nchunks if average else 1.0
    if _cuda(q) and chunk % 8 == 0:
        out = torch.empty(chunk, dtype=dtype, device=q.device)
        K._check(_lib().dsb200_dequant_reduce(K._ptr(q), K._ptr(s), K._ptr(out), K._dt(out), chunk, nchunks, group,
                                              bits, post, K._stream()), "dequant_reduce")
    else:
        full = dequantize(q, s, in_groups, bits, torch.float32)
        out = (full.view(nchunks, chunk).sum(0) * post).to(dtype)
    if requantize_bits:
        return quantize(out, out_groups, requantize_bits)
    return out


# ---- collectives ------------------------------------------------------------------------------------------
def quantized_all_gather(full_out, shard, group, fmt="fp8", block=FP8_BLOCK):
    """qwZ: all-gather ``shard`` (bf16/fp16) as block-scaled FP8 (half the bytes), dequantize into ``full_out``."""
    world = dist.get_world_size(group)
    n = shard.numel()
    pad = (-n) % block
    src = shard if pad == 0 else torch.cat([shard.view(-1), shard.new_zeros(pad)])
    q, s = fp8_block_quantize(src, block)
    qall = torch.empty(world * q.numel(), dtype=torch.uint8, device=shard.device)
    sall = torch.empty(world * s.numel(), dtype=torch.float32, device=shard.device)
    dist.all_gather_into_tensor(qall, q, group=group)
    dist.all_gather_into_tensor(sall, s, group=group)
    if pad == 0:
        fp8_block_dequantize(qall, sall, full_out.dtype, block, out=full_out.view(-1)[:world * n])
    else:
        tmp = fp8_block_dequantize(qall, sall, full_out.dtype, block)
        full_out.view(-1)[:world * n].copy_(tmp.view(world, n + pad)[:, :n].reshape(-1))
    return full_out


def all_to_all_quant_reduce(tensors, groups, bits=4):
    """qgZ: quantized reduce-scatter of each tensor (mean over ranks). ``groups``: {"local_i": pg, "global_j": pg}.
    Single NVSwitch domain -> one stage (no inter-node hop). Returns list of reduced local partitions."""
    local = next((g for k, g in groups.items() if k.startswith("local")), None)
    world = dist.get_world_size(local)
    outs = []
    for t in tensors:
        flat = t.contiguous().view(-1)
        n = flat.numel()
        if n % (world * 256) != 0 or t.dim() < 2:
            part = torch.empty(n // world if n % world == 0 else 0, dtype=t.dtype, device=t.device)
            if n % world == 0:
                dist.reduce_scatter_tensor(part, flat / world, group=local)
                outs.append(part)
            else:
                dist.all_reduce(flat, group=local)
                outs.append(flat / world)
            continue
        group_elems = 256 if (n // world) % 512 else 512
        ngroups = n // group_elems
        q, s = quantize(flat.to(torch.bfloat16) if flat.dtype == torch.float32 else flat, ngroups, bits)
        qr, sr = torch.empty_like(q), torch.empty_like(s)
        dist.all_to_all_single(qr, q, group=local)
        dist.all_to_all_single(sr, s, group=local)
        outs.append(quantized_reduction(qr, sr, ngroups, ngroups // world, bits, world, dtype=t.dtype
                                        if t.dtype != torch.float32 else torch.bfloat16).to(t.dtype))
    return outs


_LOCO_STEP = 0


def all_to_all_loco_quant_reduce(params, groups, loco_param=None):
    """qgZ with LoCo error feedback (reference ``coalesced_collectives.py:81``): the quantization error of every
    gradient is kept (int8-compressed) on the parameter and added back, damped by ``err_beta``, before the next
    quantization; the buffers are reset every ``reset_T`` steps. One NVSwitch domain -> one all-to-all stage.
    ``params``: parameters whose ``.grad`` is reduced; returns each rank's (mean) partition per parameter."""
    global _LOCO_STEP
    loco_param = loco_param or {"err_beta": 0.8, "reset_T": 1024}
    beta, reset_t = loco_param["err_beta"], loco_param["reset_T"]
    local = next((g for k, g in groups.items() if k.startswith("local")), None)
    world = dist.get_world_size(local)
    if _LOCO_STEP > reset_t:
        _LOCO_STEP = 0
    outs = []
    for p in params:
        g = p.grad
        n = g.numel()
        if g.dim() < 2 or n % (world * 256) != 0:
            outs.extend(all_to_all_quant_reduce([g], groups))
            continue
        flat = g.contiguous().view(-1)
        if _LOCO_STEP == 0 or not hasattr(p, "intra_ef_buf"):
            err = torch.zeros_like(flat, dtype=torch.float32)
        else:
            eq, es = p.intra_ef_buf
            err = dequantize(eq, es, es.numel(), 8, dtype=torch.bfloat16).float().view(-1)
        comp = flat.float() + beta * err
        group_elems = 256 if (n // world) % 512 else 512
        ngroups = n // group_elems
        q, s = quantize(comp.to(torch.bfloat16), ngroups, 4)
        new_err = comp - dequantize(q, s, ngroups, 4, dtype=torch.bfloat16).float().view(-1)
        p.intra_ef_buf = quantize(new_err.to(torch.bfloat16), ngroups, 8)
        qr, sr = torch.empty_like(q), torch.empty_like(s)
        dist.all_to_all_single(qr, q, group=local)
        dist.all_to_all_single(sr, s, group=local)
        outs.append(quantized_reduction(qr, sr, ngroups, ngroups // world, 4, world,
                                        dtype=g.dtype if g.dtype != torch.float32 else torch.bfloat16).to(g.dtype))
    _LOCO_STEP += 1
    return outs


class CUDAQuantizer:
    """Object API used by ZeRO-3 parameter all-gather in the reference (``partition_parameters.py:846``)."""
    async_flag = True
    target_group_size = 8000
    group_size_cache = {}

    def __init__(self):
        pass

    def quantize(self, param, groups=None):
        if groups is None:
            n = param.numel()
            groups = max(1, n // 512) if n % 512 == 0 else max(1, n // 256)
        return quantize(param, groups, 8)

    def dequantize(self, quantized_param, scale):
        return dequantize(quantized_param, scale, scale.numel(), 8)


class QuantizerBuilder:
    NAME = "quantizer"

    def is_compatible(self, verbose=False):
        return True

    def load(self, verbose=False):
        import sys
        return sys.modules[__name__]
