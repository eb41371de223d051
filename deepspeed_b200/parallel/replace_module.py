"""Layer replacement for the v1 inference / BERT-era training path (reference ``module_inject/replace_module.py``,
``replace_policy.py``, ``module_quantize.py``, ``auto_tp.py::ReplaceWithTensorSlicing``).

``replace_transformer_layer`` swaps every HuggingFace BERT-style encoder layer (``attention.self.{query,key,value}``,
``attention.output``, ``intermediate``, ``output``) for :class:`DeepSpeedTransformerLayer` (fused QKV projection, fused
norm kernels), taking over the weights; ``revert_transformer_layer`` is the inverse. Decoder families are handled by
AutoTP plus the fused ops of ``deepspeed_b200.models`` rather than by per-architecture containers."""
import torch
import torch.nn as nn

from .. import comm as dist
from ..utils.logging import log_dist


class DSPolicy:
    """Base of the weight-extraction policies (reference ``module_inject/policy.py``)."""
    _orig_layer_class = None
    cuda_graph_supported = False

    def __init__(self, inference=True, linear_layer=True, scale_attention=True, pre_attn_norm=False):
        self.inference, self.linear_layer = inference, linear_layer
        self.scale_attention, self.pre_attn_norm = scale_attention, pre_attn_norm

    def attention(self):
        raise NotImplementedError

    def get_hidden_heads(self):
        raise NotImplementedError

    def mlp(self):
        raise NotImplementedError

    def layernorm(self):
        raise NotImplementedError


class HFBertLayerPolicy(DSPolicy):
    """Weights of a HuggingFace ``BertLayer`` (post-LN) or a pre-LN variant with ``PreAttentionLayerNorm`` /
    ``PostAttentionLayerNorm`` / ``intermediate.dense_act``."""

    def __init__(self, client_module, inference=False):
        super().__init__(inference, pre_attn_norm=hasattr(client_module, "PreAttentionLayerNorm"))
        self.client_module = client_module
        self.cuda_graph_supported = True
        if HFBertLayerPolicy._orig_layer_class is None:
            try:
                import transformers
                HFBertLayerPolicy._orig_layer_class = [transformers.models.bert.modeling_bert.BertLayer,
                                                       transformers.models.roberta.modeling_roberta.RobertaLayer]
            except Exception:  # noqa: BLE001
                HFBertLayerPolicy._orig_layer_class = None

    def get_hidden_heads(self):
        a = self.client_module.attention
        eps = (self.client_module.PostAttentionLayerNorm if self.pre_attn_norm else a.output.LayerNorm).eps
        return a.self.query.weight.shape[1], a.self.num_attention_heads, eps

    def attention(self):
        s, o = self.client_module.attention.self, self.client_module.attention.output.dense
        return (s.query, s.key, s.value), o

    def mlp(self):
        inter = self.client_module.intermediate
        return (inter.dense_act if self.pre_attn_norm else inter.dense), self.client_module.output.dense

    def layernorm(self):
        m = self.client_module
        if self.pre_attn_norm:
            return m.PostAttentionLayerNorm, m.PreAttentionLayerNorm
        return m.attention.output.LayerNorm, m.output.LayerNorm


def _walk_replace(model, match, build):
    """Replace every sub-module for which ``match(m)`` holds by ``build(m, index)``; returns the count."""
    n = 0
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if match(child):
                setattr(parent, name, build(child, n))
                n += 1
    return n


def replace_transformer_layer(orig_layer_impl, model, checkpoint_dict=None, config=None, model_config=None,
                              training=False):
    """Replace BERT-style layers of class ``orig_layer_impl`` with the fused transformer layer. ``config``: inference
    config (``dtype``) or None; ``model_config``: the HF config (sizes, dropout, eps). Returns ``model``."""
    from ..ops.transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer
    classes = tuple(orig_layer_impl) if isinstance(orig_layer_impl, (list, tuple)) else (orig_layer_impl, )
    dtype = getattr(config, "dtype", None)

    def build(child, idx):
        pol = HFBertLayerPolicy(child, inference=not training)
        hidden, heads, eps = pol.get_hidden_heads()
        (q, k, v), attn_out = pol.attention()
        inter, out = pol.mlp()
        attn_ln, out_ln = pol.layernorm()
        mc = model_config
        cfg = DeepSpeedTransformerConfig(
            batch_size=-1, hidden_size=hidden, intermediate_size=inter.weight.shape[0], heads=heads,
            attn_dropout_ratio=getattr(mc, "attention_probs_dropout_prob", 0.0) if training else 0.0,
            hidden_dropout_ratio=getattr(mc, "hidden_dropout_prob", 0.0) if training else 0.0,
            num_hidden_layers=getattr(mc, "num_hidden_layers", -1), initializer_range=getattr(
                mc, "initializer_range", 0.02), layer_norm_eps=eps, pre_layer_norm=pol.pre_attn_norm,
            fp16=dtype == torch.float16, training=training)
        ws = [q.weight, k.weight, v.weight, attn_out.weight, attn_ln.weight, inter.weight, out.weight, out_ln.weight]
        bs = [q.bias, k.bias, v.bias, attn_out.bias, attn_ln.bias, inter.bias, out.bias, out_ln.bias]
        layer = DeepSpeedTransformerLayer(cfg, initial_weights=ws, initial_biases=bs)
        layer._replaced_from = type(child)
        if dtype is not None and dtype.is_floating_point:
            layer.to(dtype)
        return layer.to(q.weight.device)

    n = _walk_replace(model, lambda m: isinstance(m, classes), build)
    log_dist(f"replace_transformer_layer: {n} layers of {[c.__name__ for c in classes]} replaced", [0])
    return model


def revert_transformer_layer(orig_layer_impl, model, config, preln=False):
    """Inverse of :func:`replace_transformer_layer`: rebuild ``orig_layer_impl(config)`` layers from the fused ones."""
    from ..ops.transformer import DeepSpeedTransformerLayer

    def build(child, idx):
        orig = orig_layer_impl(config)
        pol = HFBertLayerPolicy(orig)
        (q, k, v), attn_out = pol.attention()
        inter, out = pol.mlp()
        attn_ln, out_ln = pol.layernorm()
        qw, kw, vw = torch.chunk(child.attn_qkvw.data, 3, dim=0)
        qb, kb, vb = torch.chunk(child.attn_qkvb.data, 3, dim=0)
        pairs = [(q.weight, qw), (q.bias, qb), (k.weight, kw), (k.bias, kb), (v.weight, vw), (v.bias, vb),
                 (attn_out.weight, child.attn_ow.data), (attn_out.bias, child.attn_ob.data),
                 (attn_ln.weight, child.attn_nw.data), (attn_ln.bias, child.attn_nb.data),
                 (inter.weight, child.inter_w.data), (inter.bias, child.inter_b.data),
                 (out.weight, child.output_w.data), (out.bias, child.output_b.data),
                 (out_ln.weight, child.norm_w.data), (out_ln.bias, child.norm_b.data)]
        for dst, src in pairs:
            dst.data = src.clone().to(dst.dtype)
        return orig.to(child.attn_qkvw.device)

    _walk_replace(model, lambda m: isinstance(m, DeepSpeedTransformerLayer), build)
    return model


def quantize_transformer_layer(orig_layer_impl, model, megatron=False, preln=False):
    """Cast the GEMM weights of every ``orig_layer_impl`` layer to int8 storage (reference ``module_quantize.py``; the
    values must already be quantized, e.g. by MoQ training)."""

    def q(linear):
        linear.weight.data = linear.weight.data.to(torch.int8)

    for m in model.modules():
        if not isinstance(m, orig_layer_impl):
            continue
        if megatron:
            for lin in (m.attention.query_key_value, m.attention.dense, m.mlp.dense_h_to_4h, m.mlp.dense_4h_to_h):
                q(lin)
        else:
            s = m.attention.self
            for lin in (s.query, s.key, s.value, m.attention.output.dense,
                        m.intermediate.dense_act if preln else m.intermediate.dense, m.output.dense):
                q(lin)
    return model


class ReplaceWithTensorSlicing:
    """Copies (slices of) full checkpoint tensors into tensor-parallel destination tensors. ``copy`` slices whichever
    of the two dims differs; ``strided_copy`` does so for a fused QKV tensor of ``num_splits`` stacked parts so that every
    rank gets matching slices of each part."""

    def __init__(self, mp_group=None, mp_size=1, out_dim=1, in_dim=0):
        self.gpu_index = dist.get_rank(group=mp_group) if mp_group is not None else 0
        self.mp_size, self.out_dim, self.in_dim = mp_size, out_dim, in_dim

    @staticmethod
    def _wrap(dst, src):
        out = nn.Parameter(dst, requires_grad=False)
        if hasattr(src, "scale"):
            out.scale = src.scale
        return out

    def _slice(self, src, dim, size):
        r = self.gpu_index
        assert src.shape[dim] >= size, "merging tensors is not supported here: merge the checkpoint first"
        return src.narrow(dim, r * size, size)

    def copy(self, dst, src, int8=False, allocate_tensor=False):
        if src is None:
            return src
        assert not dst.is_meta, "destination must be materialised (copy_ silently does nothing on meta tensors)"
        if allocate_tensor:
            dst = torch.empty_like(dst)
        if src.dim() == 2 and dst.dim() == 2:
            piece = src
            for d in (0, 1):
                if src.shape[d] != dst.shape[d]:
                    piece = self._slice(piece, d, dst.shape[d])
            dst.data.copy_(piece)
        elif src.shape[0] == dst.shape[0]:
            dst.data.copy_(src)
        else:
            dst.data.copy_(self._slice(src, 0, dst.shape[0]))
        return self._wrap(dst.data, src)

    def strided_copy(self, dst, src, num_splits, int8=False, allocate_tensor=False):
        if src is None:
            return src
        if allocate_tensor:
            dst = torch.empty_like(dst)
        fused_dim = 0 if (int8 or src.dim() == 1) else src.dim() - 1
        if src.shape[fused_dim] == dst.shape[fused_dim if src.dim() == 1 else self.out_dim if src.dim() == 2 else 0]:
            dst.data.reshape(-1).copy_(src.data.reshape(-1))
            return self._wrap(dst.data.reshape(src.shape), src)
        parts = torch.chunk(src.data, num_splits, dim=fused_dim)
        per = parts[0].shape[fused_dim] // self.mp_size
        mine = torch.cat([self._slice(p, fused_dim, per) for p in parts], dim=fused_dim)
        dst.data.reshape(-1).copy_(mine.reshape(-1))
        return self._wrap(dst.data.reshape(mine.shape), src)


class GroupQuantizer:
    """Symmetric int8 group quantizer used when injecting int8 inference weights (reference ``replace_module.py:44``):
    returns an int8 Parameter with ``.scale`` (per group, plus the two half-tensor scales for 2-way TP)."""

    def __init__(self, q_int8=True, group_size=1, num_bits=8, num_groups=0):
        self.q_int8, self.group_size, self.num_bits, self.num_groups = q_int8, group_size, num_bits, num_groups

    def quantize(self, inputs, qkv=True, count=1, parallel_dim=0):
        if not self.q_int8 or not qkv:
            out = nn.Parameter(inputs, requires_grad=False)
            out.scale = torch.empty(1)
            return out
        q_range = 2**self.num_bits
        groups = self.num_groups if self.num_groups > 0 else inputs.shape[0] // self.group_size

        def scales(t):
            flat = t.reshape(groups, -1).float()
            return (flat.abs().amax(dim=1, keepdim=True) * 2.0 / q_range).clamp(min=1e-12)

        s = scales(inputs)
        q = (inputs.reshape(groups, -1).float() / s).round().clamp(-q_range // 2, q_range // 2 - 1)
        out = nn.Parameter(q.reshape(inputs.shape).to(torch.int8).contiguous(), requires_grad=False)
        halves = inputs.split(inputs.shape[parallel_dim] // 2, dim=parallel_dim)
        out.scale = torch.cat([s.squeeze().unsqueeze(0)] + [scales(h).squeeze().unsqueeze(0) for h in halves[:2]],
                              dim=0).reshape(groups, -1).contiguous()
        return out


def generic_injection(module, dtype=None, enable_cuda_graph=True):
    """Reference ``module_inject/replace_module.py::generic_injection``: in a diffusers pipeline (or any module tree)
    replace every attention module that looks like diffusers' ``Attention`` / ``CrossAttention`` (``to_q`` / ``to_k`` /
    ``to_v`` / ``to_out``) by :class:`DeepSpeedDiffusersAttention` and every ``BasicTransformerBlock`` (``norm1..3``,
    ``attn1``, ``attn2``, GEGLU ``ff``) by :class:`DeepSpeedDiffusersTransformerBlock`; the sub-models are cast to
    ``dtype``. Returns the module; ``module.ds_injected`` counts what was replaced."""
    from ..ops.transformer.inference.config import DeepSpeedInferenceConfig
    from ..ops.transformer.inference.diffusers_2d_transformer import Diffusers2DTransformerConfig
    from ..ops.transformer.inference.diffusers_attention import DeepSpeedDiffusersAttention
    from ..ops.transformer.inference.diffusers_transformer_block import DeepSpeedDiffusersTransformerBlock
    roots = [getattr(module, n) for n in ("unet", "vae", "text_encoder") if isinstance(getattr(module, n, None), nn.Module)]
    if not roots and isinstance(module, nn.Module):
        roots = [module]
    if dtype is not None:
        for r in roots:
            r.to(dtype)

    def is_attention(m):
        return all(hasattr(m, n) for n in ("to_q", "to_k", "to_v", "to_out")) and hasattr(m, "heads") and \
            isinstance(m.to_q, nn.Linear) and not isinstance(m, DeepSpeedDiffusersAttention)

    def is_block(m):
        ff = getattr(m, "ff", None)
        return all(hasattr(m, n) for n in ("norm1", "norm2", "norm3", "attn1", "attn2")) and ff is not None and \
            hasattr(ff, "net") and hasattr(ff.net[0], "proj") and getattr(m, "attn2", None) is not None and \
            not isinstance(m, DeepSpeedDiffusersTransformerBlock)

    n_attn = n_block = 0
    for root in roots:
        for parent in list(root.modules()):
            for name, child in list(parent.named_children()):
                if is_attention(child):
                    w = child.to_q.weight
                    cfg = DeepSpeedInferenceConfig(hidden_size=w.shape[0], heads=child.heads, dtype=w.dtype)
                    setattr(parent, name, DeepSpeedDiffusersAttention.from_module(child, cfg))
                    n_attn += 1
        for parent in list(root.modules()):
            for name, child in list(parent.named_children()):
                if is_block(child):
                    setattr(parent, name, DeepSpeedDiffusersTransformerBlock(child, Diffusers2DTransformerConfig()))
                    n_block += 1
    try:
        module.ds_injected = {"attention": n_attn, "transformer_block": n_block}
    except Exception:  # noqa: BLE001
        pass
    return module


# ---- generic class-keyed replacement (reference ``replace_module.py:630`` / ``module_quantize.py:57``) -----------
def _swap_children(model, policy, state_dict=None, prefix="", counter=None):
    """Depth-first: a child whose class is a key of ``policy`` is replaced by ``fn(child, policy_cls, layer_id[,
    prefix, state_dict])``; other children are searched. Returns the number of replacements."""
    counter = counter if counter is not None else [0]
    for name, child in list(model.named_children()):
        entry = policy.get(child.__class__)
        if entry is not None:
            fn, pol = entry if isinstance(entry, tuple) else (entry, None)
            try:
                new = fn(child, pol, counter[0], prefix=prefix + name, state_dict=state_dict)
            except TypeError:
                new = fn(child, pol, counter[0]) if pol is not None else fn(child)
            setattr(model, name, new)
            counter[0] += 1
        else:
            _swap_children(child, policy, state_dict, prefix + name + ".", counter)
    return counter[0]


def replace_module(model, orig_class, replace_fn, _replace_policy, checkpoint=None):
    """Replace every instance of ``orig_class`` (or, with ``orig_class=None``, of the classes the registered policies
    name) using ``replace_fn(child, policy, layer_id)``. ``checkpoint``: a ``.pt`` / ``.safetensors`` file whose state
    dict is handed to ``replace_fn`` when it accepts ``state_dict``."""
    sd = None
    if checkpoint is not None:
        if checkpoint.endswith(".safetensors"):
            from safetensors.torch import load_file
            sd = load_file(checkpoint)
        else:
            sd = torch.load(checkpoint, map_location="cpu", weights_only=False)
    policy = {}
    if orig_class is not None:
        policy[orig_class] = (replace_fn, _replace_policy)
    else:
        for pol in (HFBertLayerPolicy, ):
            cls = getattr(pol, "_orig_layer_class", None)
            for c in (cls if isinstance(cls, (list, tuple)) else [cls]):
                if c is not None:
                    policy[c] = (replace_fn, pol)
    assert policy, ("No default policy found! Please specify your policy injection_policy (like "
                    "{BertLayer: HFBertLayerPolicy}).")
    _swap_children(model, policy, state_dict=sd)
    return model


def quantize_module(model, orig_class, quantize_fn):
    """Replace every ``orig_class`` child by ``quantize_fn(child)``."""
    _swap_children(model, {orig_class: quantize_fn})
    return model


def skip_level_0_prefix(model, state_dict):
    """Bloom / OPT checkpoints name their weights without the top-level module prefix - unless the keys say
    otherwise (``model.`` prefix present)."""
    import re
    text = str(model)
    m = re.search(r": (.*?)Model", text) or re.search(r": (.*?)Stack", text) or re.match(r"(.*?)Model", text)
    if state_dict is not None and any(re.match(r"^model[.]", k) for k in state_dict):
        return False
    return m is not None and m.group(1).lower() in ("bloom", "opt")


def get_transformer_name(replaced_module):
    """Dotted path of the ``ModuleList`` holding the transformer layers (``"transformer.h"``, ``"model.layers"``)."""
    for n, c in replaced_module.named_children():
        for name, child in c.named_children():
            if isinstance(child, nn.ModuleList):
                return f"{n}.{name}"
    return ""
