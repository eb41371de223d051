"""ALiBi / attention-bias helpers that follow the head sharding of AutoTP.

Reference: ``module_inject/auto_tp_model_utils.py`` (``build_bloom_alibi_tensor``, ``build_mpt_atten_bias_tensor``,
``build_mpt_alibi_tensor``, ``get_alibi_mask``, ``install_head_sharded_helper``). Models with ALiBi build one bias
slope per attention head at the *model* level and hand the tensor to every layer; once the attention projections are
sharded a rank computes only its heads and must see only their biases.

Approach here: leave the model's own builder untouched, call it for all heads and keep this rank's head range - the
bias is ``O(heads * seq)`` and built once per forward, so recomputing slopes per rank (what the reference does by
re-implementing each family's formula) buys nothing. The wrapper is installed on the *instance* (a class-wide patch
would outlive the model and hit instances that were never sharded)."""
import functools

import torch

from .. import comm as dist
from .fusedqkv_utils import get_shard_size_list


def get_head_shard_sizes(num_heads, mp_group=None, num_kv_heads=None):
    return get_shard_size_list(num_heads, dist.get_world_size(group=mp_group), num_kv_heads=num_kv_heads)


def _head_range(num_heads, mp_group=None, head_shard_sizes=None):
    world, rank = dist.get_world_size(group=mp_group), dist.get_rank(group=mp_group)
    sizes = head_shard_sizes or get_shard_size_list(num_heads, world, num_kv_heads=num_heads)
    if len(sizes) != world or sum(sizes) != num_heads:
        raise ValueError(f"Head shard sizes {sizes} do not partition {num_heads} heads across {world} ranks")
    return sum(sizes[:rank]), sizes[rank]


def slice_heads(bias, num_heads, head_dim, mp_group=None, head_shard_sizes=None, folded_batch=False):
    """This rank's heads of an attention bias. ``head_dim``: the axis holding the heads; ``folded_batch``: the axis
    holds ``batch * heads`` (Bloom's ``[batch * heads, 1, seq]``)."""
    lo, n = _head_range(num_heads, mp_group, head_shard_sizes)
    if folded_batch:
        shape = list(bias.shape)
        b = shape[head_dim] // num_heads
        view = bias.reshape(shape[:head_dim] + [b, num_heads] + shape[head_dim + 1:])
        out = view.narrow(head_dim + 1, lo, n)
        return out.reshape(shape[:head_dim] + [b * n] + shape[head_dim + 1:])
    return bias.narrow(head_dim, lo, n)


def install_head_sharded_helper(module, name, wrapper, mp_group=None, num_heads=None, num_kv_heads=None):
    """Bind ``wrapper(module, *args, mp_group=, head_shard_sizes=, total_num_heads=, **kw)`` over ``module.<name>``;
    the original stays reachable as ``module.<name>_orig``. Installing twice keeps the first original."""
    orig = f"{name}_orig"
    if orig not in module.__dict__:
        setattr(module, orig, getattr(module, name))
    sizes = get_head_shard_sizes(num_heads, mp_group, num_kv_heads) if num_heads is not None else None
    setattr(module, name, functools.partial(wrapper, module, mp_group=mp_group, head_shard_sizes=sizes,
                                            total_num_heads=num_heads))


def build_bloom_alibi_tensor(module, attention_mask, num_heads, dtype, mp_group=None, head_shard_sizes=None,
                             total_num_heads=None):
    """Bloom / Falcon: ``[batch * heads, 1, seq]`` -> the local heads."""
    total = total_num_heads or num_heads
    full = module.build_alibi_tensor_orig(attention_mask, total, dtype)
    return slice_heads(full, total, 0, mp_group, head_shard_sizes, folded_batch=True)


def build_mpt_alibi_tensor(module, num_heads, sequence_length, alibi_bias_max=8, device=None, mp_group=None,
                           head_shard_sizes=None, total_num_heads=None):
    """MPT: ``[heads, 1, seq]`` -> the local heads."""
    total = total_num_heads or num_heads
    full = module.build_mpt_alibi_tensor_orig(total, sequence_length, alibi_bias_max, device)
    return slice_heads(full, total, 0, mp_group, head_shard_sizes)


def build_mpt_atten_bias_tensor(module, *args, mp_group=None, head_shard_sizes=None, total_num_heads=None, **kwargs):
    """Older MPT (``_attn_bias``): returns ``(attn_bias [1, heads, 1, seq], attention_mask)``."""
    attn_bias, attention_mask = module._attn_bias_orig(*args, **kwargs)
    if attn_bias is not None and total_num_heads:
        attn_bias = slice_heads(attn_bias, total_num_heads, 1, mp_group, head_shard_sizes)
    return attn_bias, attention_mask


def get_alibi_mask(module, tensor, seq_length_with_past, mp_group=None, head_shard_sizes=None, total_num_heads=None):
    """Baichuan: ``[heads, seq, seq]`` mask -> the local heads."""
    mask = module.get_alibi_mask_orig(tensor, seq_length_with_past)
    total = total_num_heads or mask.shape[0]
    return slice_heads(mask, total, 0, mp_group, head_shard_sizes)


_BUILDERS = (("build_alibi_tensor", build_bloom_alibi_tensor), ("build_mpt_alibi_tensor", build_mpt_alibi_tensor),
             ("_attn_bias", build_mpt_atten_bias_tensor), ("get_alibi_mask", get_alibi_mask))


def install_alibi_head_sharding(model, mp_group, num_heads):
    """Wrap every ALiBi builder found on the model's modules; returns how many were wrapped."""
    n = 0
    if not num_heads or dist.get_world_size(group=mp_group) == 1:
        return n
    for m in model.modules():
        for name, wrapper in _BUILDERS:
            if callable(getattr(m, name, None)) and not isinstance(getattr(m, name), torch.nn.Module):
                install_head_sharded_helper(m, name, wrapper, mp_group, num_heads, num_kv_heads=num_heads)
                n += 1
    return n
