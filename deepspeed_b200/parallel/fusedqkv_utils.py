"""Tensor-parallel sharding of fused QKV projections, by layout family.

Reference: ``module_inject/fusedqkv_utils.py`` (``prepare_tp_fused_qkvw`` and friends) and ``tp_shard.py``. Model
families lay the rows of a fused ``[q; k; v]`` weight out differently, and a rank must receive rows that form whole
heads with matching q / k / v:

* ``bloomtype`` (Bloom, Falcon, GPT-NeoX): interleaved per head ``[q1 k1 v1 q2 k2 v2 ...]`` - a contiguous row range
* ``glmtype`` / ``qwentype`` (GLM, MPT, Baichuan, QWen): ``[q(all heads); k; v]`` - this rank's range of each third
  (ChatGLM2/3 with two KV groups: ``q`` is ``n_embd`` rows, ``k`` / ``v`` the rest)
* ``phi3type``: ``[q; k; v]`` with grouped KV heads - ranges of ``q``, ``k`` and ``v`` by their own head counts
* ``bigcodetype`` (GPT-BigCode, multi-query): ``[q; kv]`` - ``q`` sharded, the single KV head replicated
* ``codegentype``: ``mp_num = 4`` blocks of ``[q; v; k]`` sub-blocks - sub-slices regrouped so that the local tensor
  is again four blocks

Everything is expressed as ONE row-index tensor per (layout, rank): the shard is ``weight[rows]`` (and ``bias[rows]``),
which also describes how to put a gathered parameter back together."""
import torch

from .tp import _shard_sizes

_STATE = {"num_kv_heads": None, "num_attention_heads": None, "n_embd": None, "tp_grain_size": 1}

FUSED_QKV_TYPE_DICT = {
    "CodeGenBlock": "codegentype",
    "BloomBlock": "bloomtype",
    "GLMBlock": "glmtype",
    "MPTBlock": "glmtype",
    "MptBlock": "glmtype",
    "BaichuanLayer": "glmtype",
    "QWenBlock": "qwentype",
    "FalconDecoderLayer": "bloomtype",
    "GPTNeoXLayer": "bloomtype",
    "GPTBigCodeBlock": "bigcodetype",
    "DecoderLayer": "glmtype",
    "Phi3DecoderLayer": "phi3type",
}
FUSED_QKV_NAMES = ("qkv_proj", "query_key_value", "attn.Wqkv", "self_attn.W_pack", "c_attn")


def set_num_kv_heads(n):
    _STATE["num_kv_heads"] = n
    from . import layers
    layers.set_num_kv_heads(n)


def set_num_attention_heads(n):
    _STATE["num_attention_heads"] = n


def set_n_embd(n):
    _STATE["n_embd"] = n


def set_tp_grain_size(n):
    _STATE["tp_grain_size"] = n


def get_num_kv_heads():
    return _STATE["num_kv_heads"]


def get_num_attention_heads():
    return _STATE["num_attention_heads"]


def get_n_embd():
    return _STATE["n_embd"]


def get_shard_size_list(total_size, mp_size, name=None, num_kv_heads=None):
    """Rows per rank. Attention projections are cut at KV-head boundaries when the head count is known (uneven head
    counts give uneven shards); everything else in multiples of the grain size, the remainder on the last rank."""
    nkv = num_kv_heads if num_kv_heads is not None else _STATE["num_kv_heads"]
    mlp_like = name is not None and any(s in str(name) for s in ("mlp", "gate_proj", "up_proj", "down_proj", "w1", "w2",
                                                                 "w3"))
    if nkv and total_size % nkv == 0 and not mlp_like and str(name) not in ("lm_head", "embed_out"):
        sizes = [total_size * (nkv // mp_size + (1 if r < nkv % mp_size else 0)) // nkv for r in range(mp_size)]
    else:
        sizes = _shard_sizes(total_size, mp_size, max(1, _STATE["tp_grain_size"])) \
            if total_size >= _STATE["tp_grain_size"] else _shard_sizes(total_size, mp_size)
    assert sum(sizes) == total_size, f"AutoTP shard sizes {sizes} for layer '{name}' do not sum to {total_size}"
    return sizes


def get_shard_size(total_size, mp_size, name=None, rank=None, mp_group=None, num_kv_heads=None):
    if rank is None:
        from .. import comm as dist
        rank = dist.get_rank(group=mp_group) if mp_group is not None else dist.get_rank()
    return get_shard_size_list(total_size, mp_size, name, num_kv_heads)[rank]


def require_tp_fused_qkvw(name, mp_size):
    return mp_size > 1 and any(n in name for n in FUSED_QKV_NAMES)


def get_fused_qkv_type(module):
    """Layout family from the class names in ``module``'s tree (the longest matching name wins: ``DecoderLayer`` is
    also a substring of ``FalconDecoderLayer``)."""
    names = {type(m).__name__ for m in module.modules()} if hasattr(module, "modules") else {str(module)}
    text = " ".join(names)
    hits = [k for k in FUSED_QKV_TYPE_DICT if k in text]
    return FUSED_QKV_TYPE_DICT[max(hits, key=len)] if hits else None


def _range_of(total, world, rank, offset=0, name=None, nkv=None):
    sizes = get_shard_size_list(total, world, name, nkv)
    lo = offset + sum(sizes[:rank])
    return torch.arange(lo, lo + sizes[rank])


def fused_qkv_row_index(fused_type, total_rows, mp_size, rank, hidden_cols=None, codegen_mp_num=4):
    """Rows of the fused weight that tensor-parallel ``rank`` keeps."""
    nkv, nh, n_embd = _STATE["num_kv_heads"], _STATE["num_attention_heads"], _STATE["n_embd"]
    if fused_type in (None, "bloomtype"):
        return _range_of(total_rows, mp_size, rank)
    if fused_type in ("glmtype", "qwentype"):
        if nkv == 2 and n_embd:
            kv = (total_rows - n_embd) // 2
            parts = [(n_embd, 0), (kv, n_embd), (kv, n_embd + kv)]
        else:
            third = total_rows // 3
            parts = [(third, 0), (third, third), (third, 2 * third)]
        return torch.cat([_range_of(n, mp_size, rank, off) for n, off in parts])
    if fused_type == "phi3type":
        assert nkv and nh, "phi3 fused QKV needs set_num_kv_heads / set_num_attention_heads"
        head_dim = (hidden_cols if hidden_cols is not None else n_embd) // nh
        q_rows = total_rows - 2 * nkv * head_dim
        parts = [(q_rows, 0), (nkv * head_dim, q_rows), (nkv * head_dim, q_rows + nkv * head_dim)]
        return torch.cat([_range_of(n, mp_size, rank, off) for n, off in parts])
    if fused_type == "bigcodetype":
        assert n_embd, "GPT-BigCode fused QKV needs set_n_embd"
        # the query heads are what gets divided; the lone KV head must not drive the split
        return torch.cat([_range_of(n_embd, mp_size, rank, nkv=nh or n_embd), torch.arange(n_embd, total_rows)])
    if fused_type == "codegentype":
        assert nkv is None or nkv % (mp_size * codegen_mp_num) == 0, \
            "codegen AutoTP requires num_kv_heads % (mp_size * mp_num) == 0"
        block, sub = total_rows // codegen_mp_num, total_rows // codegen_mp_num // 3
        piece = sub // mp_size
        rows = []
        for j in range(codegen_mp_num):  # local block j is sub-slice r of global block blk
            blk, r = divmod(rank * codegen_mp_num + j, mp_size)
            for part in range(3):
                lo = blk * block + part * sub + r * piece
                rows.append(torch.arange(lo, lo + piece))
        return torch.cat(rows)
    raise ValueError(f"unknown fused_qkv_type {fused_type!r}")


def fused_qkv_subparam_sizes(module, weight_shape):
    """Sizes of the contiguous sub-parameters the fused weight is cut into (what a gather / checkpoint conversion needs
    to undo the split), or ``None`` for the layouts that interleave or replicate blocks."""
    t = get_fused_qkv_type(module) or "bloomtype"
    total = weight_shape[0]
    if t == "bloomtype":
        return (total, )
    if t in ("glmtype", "qwentype"):
        if _STATE["num_kv_heads"] == 2 and _STATE["n_embd"]:
            kv = (total - _STATE["n_embd"]) // 2
            return (_STATE["n_embd"], kv, kv)
        return (total // 3, ) * 3
    if t == "phi3type" and _STATE["num_kv_heads"] and _STATE["num_attention_heads"] and len(weight_shape) > 1:
        hd = weight_shape[1] // _STATE["num_attention_heads"]
        kv = _STATE["num_kv_heads"] * hd
        return (total - 2 * kv, kv, kv)
    return None


def prepare_tp_fused_qkvw(module, src, mp_size, gpu_index):
    """This rank's rows of the fused weight (or bias) ``src`` of ``module``'s block."""
    if src is None:
        return None
    t = get_fused_qkv_type(module)
    if t is None:
        from ..utils.logging import logger
        logger.warning("Unrecognized fused QKV layout: sharding as contiguous per-head rows (bloom layout)")
    cols = src.shape[1] if src.dim() > 1 else None
    return src[fused_qkv_row_index(t, src.shape[0], mp_size, gpu_index, cols).to(src.device)]


def split_by_qkvlist_and_refuse(qkv_list, split_size, split_dim=0, cat_dim=0):
    """Split each of q / k / v the same way and fuse the i-th pieces back together: one fused tensor per rank."""
    pieces = [torch.split(m, split_size, dim=split_dim) for m in qkv_list]
    return [torch.cat([p[i] for p in pieces], dim=cat_dim) for i in range(len(pieces[0]))]
