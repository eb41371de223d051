"""HuggingFace config / checkpoint -> :class:`GenericDecoder` for every FastGen family (reference: one
``*_policy.py`` + ``*_containers.py`` pair per family under ``inference/v2/model_implementations``)."""
import torch

from .arch import ArchSpec, GenericDecoder


def _theta(cfg, default=10000.0):
    rp = getattr(cfg, "rope_parameters", None)
    if isinstance(rp, dict) and "rope_theta" in rp:
        return float(rp["rope_theta"])
    return float(getattr(cfg, "rope_theta", default) or default)


def _common(cfg, **kw):
    base = dict(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
                num_attention_heads=cfg.num_attention_heads,
                max_position_embeddings=getattr(cfg, "max_position_embeddings", 4096),
                tie_word_embeddings=bool(getattr(cfg, "tie_word_embeddings", False)))
    base.update(kw)
    return base


def _llama_like(family, **over):

    def spec(cfg, dtype):
        kw = _common(cfg, family=family, num_key_value_heads=getattr(cfg, "num_key_value_heads", None)
                     or cfg.num_attention_heads, intermediate_size=cfg.intermediate_size,
                     head_dim=getattr(cfg, "head_dim", None), norm="rms", norm_eps=cfg.rms_norm_eps, act="silu",
                     gated_mlp=True, rope_theta=_theta(cfg), dtype=dtype)
        kw.update(over)
        return ArchSpec(**kw)

    return spec


def _spec_moe(family, shared=False, **over):
    base = _llama_like(family, **over)

    def spec(cfg, dtype):
        s = base(cfg, dtype)
        s.num_experts = getattr(cfg, "num_local_experts", None) or getattr(cfg, "num_experts")
        s.num_experts_per_tok = cfg.num_experts_per_tok
        s.norm_topk_prob = bool(getattr(cfg, "norm_topk_prob", True))
        s.moe_intermediate_size = getattr(cfg, "moe_intermediate_size", None) or cfg.intermediate_size
        if shared:
            s.shared_expert_intermediate_size = cfg.shared_expert_intermediate_size
        return s

    return spec


def _spec_phi(cfg, dtype):
    hd = cfg.hidden_size // cfg.num_attention_heads
    return ArchSpec(**_common(cfg, family="phi", num_key_value_heads=getattr(cfg, "num_key_value_heads", None)
                              or cfg.num_attention_heads, intermediate_size=cfg.intermediate_size, norm="ln",
                              norm_eps=cfg.layer_norm_eps, act="gelu_tanh", gated_mlp=False, qkv_bias=True, o_bias=True,
                              mlp_bias=True, rotary_dim=int(hd * getattr(cfg, "partial_rotary_factor", 0.5)),
                              rope_theta=_theta(cfg), parallel_block=True, lm_head_bias=True, dtype=dtype))


def _spec_falcon(cfg, dtype):
    new = bool(getattr(cfg, "new_decoder_architecture", False))
    nkv = cfg.num_kv_heads if new else (1 if getattr(cfg, "multi_query", True) else cfg.num_attention_heads)
    assert not getattr(cfg, "alibi", False), "falcon with alibi positions is not supported"
    return ArchSpec(**_common(cfg, family="falcon", num_key_value_heads=nkv,
                              intermediate_size=getattr(cfg, "ffn_hidden_size", None) or 4 * cfg.hidden_size,
                              norm="ln", norm_eps=cfg.layer_norm_epsilon, act="gelu", gated_mlp=False,
                              qkv_bias=bool(cfg.bias), o_bias=bool(cfg.bias), mlp_bias=bool(cfg.bias),
                              rope_theta=_theta(cfg), parallel_block=bool(cfg.parallel_attn),
                              parallel_two_norms=new and getattr(cfg, "num_ln_in_parallel_attn", 2) == 2,
                              dtype=dtype))


def _spec_opt(cfg, dtype):
    assert cfg.do_layer_norm_before and cfg.word_embed_proj_dim == cfg.hidden_size, "only pre-LN OPT variants"
    return ArchSpec(**_common(cfg, family="opt", num_key_value_heads=cfg.num_attention_heads,
                              intermediate_size=cfg.ffn_dim, norm="ln", norm_eps=1e-5, act=cfg.activation_function,
                              gated_mlp=False, qkv_bias=cfg.enable_bias, o_bias=cfg.enable_bias,
                              mlp_bias=cfg.enable_bias, pos="learned", pos_offset=2, tie_word_embeddings=True,
                              dtype=dtype))


def _spec_qwen1(cfg, dtype):
    return ArchSpec(vocab_size=cfg.vocab_size, hidden_size=cfg.hidden_size, num_hidden_layers=cfg.num_hidden_layers,
                    num_attention_heads=cfg.num_attention_heads, num_key_value_heads=cfg.num_attention_heads,
                    intermediate_size=cfg.intermediate_size // 2, norm="rms", norm_eps=cfg.layer_norm_epsilon,
                    act="silu", gated_mlp=True, qkv_bias=True, rope_theta=float(getattr(cfg, "rotary_emb_base", 1e4)),
                    max_position_embeddings=getattr(cfg, "seq_length", 8192), family="qwen", dtype=dtype)


SPECS = {
    "llama": _llama_like("llama"),
    "mistral": _llama_like("mistral"),
    "qwen2": _llama_like("qwen2", qkv_bias=True),
    "phi3": _llama_like("phi3"),
    "exaone4": _llama_like("exaone4", qk_norm=True, sublayer_post_norm=True),
    "mixtral": _spec_moe("mixtral"),
    "qwen2_moe": _spec_moe("qwen2_moe", shared=True, qkv_bias=True),
    "phi": _spec_phi,
    "falcon": _spec_falcon,
    "opt": _spec_opt,
    "qwen": _spec_qwen1,
}
# reference directory name -> HF model_type
REFERENCE_FAMILIES = {"llama_v2": "llama", "mistral": "mistral", "mixtral": "mixtral", "opt": "opt", "falcon": "falcon",
                      "phi": "phi", "phi3": "phi3", "qwen": "qwen", "qwen_v2": "qwen2", "qwen_v2_moe": "qwen2_moe",
                      "exaone4": "exaone4", "exaone4_5": "exaone4"}


# ---------------------------------------------------------------------------------------------------------------
# checkpoint loading
# ---------------------------------------------------------------------------------------------------------------
def _first(sd, *names):
    for n in names:
        if n in sd:
            return sd[n]
    return None


def _falcon_qkv(w, spec, new_arch, multi_query):
    """Falcon's fused QKV layouts -> [q | k | v] rows. ``w``: [rows, hidden] or [rows] (bias)."""
    nh, nkv, hd = spec.num_attention_heads, spec.num_key_value_heads, spec.head_dim
    tail = w.shape[1:]
    if new_arch:    # [kv_groups, q_per_group + 2, hd]
        g = w.view(nkv, nh // nkv + 2, hd, *tail)
        q, k, v = g[:, :-2].reshape(nh * hd, *tail), g[:, -2].reshape(nkv * hd, *tail), g[:, -1].reshape(nkv * hd, *tail)
    elif multi_query:  # [nh + 2, hd]
        g = w.view(nh + 2, hd, *tail)
        q, k, v = g[:-2].reshape(nh * hd, *tail), g[-2].reshape(hd, *tail), g[-1].reshape(hd, *tail)
    else:           # [nh, 3, hd]
        g = w.view(nh, 3, hd, *tail)
        q, k, v = (g[:, i].reshape(nh * hd, *tail) for i in range(3))
    return torch.cat([q, k, v], dim=0)


@torch.no_grad()
def load_hf_weights(model: GenericDecoder, sd, hf_config=None):
    """Copy a HuggingFace state dict (any of the supported families) into ``model``."""
    spec = model.spec
    sd = dict(sd)

    def put(param, value):
        assert value is not None, f"missing tensor for {tuple(param.shape)}"
        assert tuple(param.shape) == tuple(value.shape), (tuple(param.shape), tuple(value.shape))
        param.copy_(value.to(param.dtype))

    put(model.embed_tokens.weight, _first(sd, "model.embed_tokens.weight", "transformer.word_embeddings.weight",
                                          "model.decoder.embed_tokens.weight", "transformer.wte.weight"))
    if spec.pos == "learned":
        put(model.embed_positions.weight, sd["model.decoder.embed_positions.weight"])
    fam = spec.family
    for i, layer in enumerate(model.layers):
        pre = next(p for p in (f"model.layers.{i}.", f"transformer.h.{i}.", f"model.decoder.layers.{i}.")
                   if any(k.startswith(p) for k in sd))
        g = lambda *names: _first(sd, *[pre + n for n in names])  # noqa: E731
        a = layer.self_attn
        # ---- attention
        for kind, dst in (("weight", a.qkv_proj.weight), ("bias", a.qkv_proj.bias)):
            if dst is None:
                continue
            q = g(f"self_attn.q_proj.{kind}")
            if q is not None:
                val = torch.cat([q, g(f"self_attn.k_proj.{kind}"), g(f"self_attn.v_proj.{kind}")], dim=0)
            elif fam == "falcon":
                val = _falcon_qkv(g(f"self_attention.query_key_value.{kind}"), spec,
                                  bool(getattr(hf_config, "new_decoder_architecture", False)),
                                  bool(getattr(hf_config, "multi_query", True)))
            else:  # phi3 `qkv_proj`, qwen `attn.c_attn`: already [q | k | v]
                val = g(f"self_attn.qkv_proj.{kind}", f"attn.c_attn.{kind}")
            put(dst, val)
        put(a.o_proj.weight, g("self_attn.o_proj.weight", "self_attn.dense.weight", "self_attn.out_proj.weight",
                               "self_attention.dense.weight", "attn.c_proj.weight"))
        if a.o_proj.bias is not None:
            put(a.o_proj.bias, g("self_attn.o_proj.bias", "self_attn.dense.bias", "self_attn.out_proj.bias",
                                 "self_attention.dense.bias"))
        if spec.qk_norm:
            put(a.q_norm, g("self_attn.q_norm.weight"))
            put(a.k_norm, g("self_attn.k_norm.weight"))
        # ---- norms
        if not spec.sublayer_post_norm:
            n1 = layer.input_layernorm
            put(n1.weight, g("input_layernorm.weight", "self_attn_layer_norm.weight", "ln_attn.weight", "ln_1.weight"))
            if n1.bias is not None:
                put(n1.bias, g("input_layernorm.bias", "self_attn_layer_norm.bias", "ln_attn.bias"))
        if hasattr(layer, "post_attention_layernorm"):
            n2 = layer.post_attention_layernorm
            put(n2.weight, g("post_attention_layernorm.weight", "final_layer_norm.weight", "ln_mlp.weight",
                             "ln_2.weight"))
            if n2.bias is not None:
                put(n2.bias, g("post_attention_layernorm.bias", "final_layer_norm.bias", "ln_mlp.bias"))
        if spec.sublayer_post_norm:
            put(layer.post_feedforward_layernorm.weight, g("post_feedforward_layernorm.weight"))
        # ---- MLP / MoE
        if spec.num_experts:
            moe = layer.mlp
            put(moe.gate.weight, g("mlp.gate.weight", "block_sparse_moe.gate.weight"))
            fused_gu, fused_dn = g("mlp.experts.gate_up_proj"), g("mlp.experts.down_proj")
            for e, ex in enumerate(moe.experts):
                if fused_gu is not None:   # transformers >= 5: [E, 2I, H] and [E, H, I]
                    put(ex.up.weight, fused_gu[e])
                    put(ex.down.weight, fused_dn[e])
                else:
                    w1 = g(f"block_sparse_moe.experts.{e}.w1.weight", f"mlp.experts.{e}.gate_proj.weight")
                    w3 = g(f"block_sparse_moe.experts.{e}.w3.weight", f"mlp.experts.{e}.up_proj.weight")
                    w2 = g(f"block_sparse_moe.experts.{e}.w2.weight", f"mlp.experts.{e}.down_proj.weight")
                    put(ex.up.weight, torch.cat([w1, w3], dim=0))
                    put(ex.down.weight, w2)
            if moe.shared is not None:
                put(moe.shared.up.weight, torch.cat([g("mlp.shared_expert.gate_proj.weight"),
                                                     g("mlp.shared_expert.up_proj.weight")], dim=0))
                put(moe.shared.down.weight, g("mlp.shared_expert.down_proj.weight"))
                put(moe.shared_gate.weight, g("mlp.shared_expert_gate.weight"))
        else:
            mlp = layer.mlp
            if spec.gated_mlp:
                gu = g("mlp.gate_up_proj.weight")
                if gu is None:
                    gate = g("mlp.gate_proj.weight", "mlp.w2.weight")
                    up = g("mlp.up_proj.weight", "mlp.w1.weight")
                    gu = torch.cat([gate, up], dim=0)
                put(mlp.up.weight, gu)
                put(mlp.down.weight, g("mlp.down_proj.weight", "mlp.c_proj.weight"))
            else:
                put(mlp.up.weight, g("mlp.fc1.weight", "fc1.weight", "mlp.dense_h_to_4h.weight"))
                put(mlp.down.weight, g("mlp.fc2.weight", "fc2.weight", "mlp.dense_4h_to_h.weight"))
                if mlp.up.bias is not None:
                    put(mlp.up.bias, g("mlp.fc1.bias", "fc1.bias", "mlp.dense_h_to_4h.bias"))
                    put(mlp.down.bias, g("mlp.fc2.bias", "fc2.bias", "mlp.dense_4h_to_h.bias"))
    if model.norm is not None:
        put(model.norm.weight, _first(sd, "model.norm.weight", "model.final_layernorm.weight", "transformer.ln_f.weight",
                                      "model.decoder.final_layer_norm.weight"))
        if model.norm.bias is not None:
            put(model.norm.bias, _first(sd, "model.final_layernorm.bias", "transformer.ln_f.bias",
                                        "model.decoder.final_layer_norm.bias"))
    if not spec.tie_word_embeddings:
        put(model.lm_head.weight, _first(sd, "lm_head.weight", "model.embed_tokens.weight"))
    if model.lm_head.bias is not None:
        put(model.lm_head.bias, sd["lm_head.bias"])
    return model


def build_hf_model(hf_model_or_config, state_dict=None, dtype=torch.bfloat16, device=None):
    """``GenericDecoder`` for a HuggingFace causal LM (pass the model, or its config + state dict)."""
    if hasattr(hf_model_or_config, "state_dict") and state_dict is None:
        cfg, state_dict = hf_model_or_config.config, hf_model_or_config.state_dict()
    else:
        cfg = hf_model_or_config
    mt = cfg.model_type
    if mt not in SPECS:
        raise ValueError(f"model_type '{mt}' has no FastGen policy here; supported: {sorted(SPECS)}")
    if mt == "exaone4" and getattr(cfg, "sliding_window", None):
        raise NotImplementedError("exaone4 hybrid (sliding window + NoPE global layers) variants are not supported")
    spec = SPECS[mt](cfg, dtype)
    model = GenericDecoder(spec, device=device)
    if state_dict is not None:
        load_hf_weights(model, state_dict, cfg)
    return model.eval()
