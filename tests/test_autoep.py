"""AutoEP: presets / config validation, router, HF block replacement (Mixtral, Qwen2/3-MoE, DeepSeek-V2/V3), expert
parallel and AutoEP folded over tensor parallelism. Mirrors the reference's tests/unit/v1/moe/test_autoep*.py."""
import copy

import pytest
import torch

from common import run_distributed
from deepspeed_b200.parallel import auto_ep_folding as F_
from deepspeed_b200.parallel import auto_ep_presets as P
from deepspeed_b200.parallel.auto_ep import (AutoEP, AutoEPMoELayer, TokenChoiceTopKRouter, apply_autoep)

TINY = dict(vocab_size=100, hidden_size=32, intermediate_size=48, num_hidden_layers=2, num_attention_heads=4,
            num_key_value_heads=2, max_position_embeddings=64)
MLA = dict(q_lora_rank=16, kv_lora_rank=16, qk_rope_head_dim=8, qk_nope_head_dim=8, v_head_dim=8)


def _hf(family):
    import transformers as T
    if family == "mixtral":
        return T.MixtralForCausalLM(T.MixtralConfig(**TINY, num_local_experts=4, num_experts_per_tok=2))
    if family == "qwen2_moe":
        return T.Qwen2MoeForCausalLM(
            T.Qwen2MoeConfig(**TINY, num_experts=4, num_experts_per_tok=2, moe_intermediate_size=24,
                             shared_expert_intermediate_size=40, decoder_sparse_step=1, norm_topk_prob=False))
    if family == "qwen3_moe":
        return T.Qwen3MoeForCausalLM(
            T.Qwen3MoeConfig(**TINY, num_experts=4, num_experts_per_tok=2, moe_intermediate_size=24,
                             decoder_sparse_step=1, norm_topk_prob=True, head_dim=8))
    if family == "deepseek_v3":
        return T.DeepseekV3ForCausalLM(
            T.DeepseekV3Config(**TINY, **MLA, n_routed_experts=8, num_experts_per_tok=3, moe_intermediate_size=24,
                               n_shared_experts=1, n_group=4, topk_group=2, first_k_dense_replace=1,
                               routed_scaling_factor=2.5, norm_topk_prob=True))
    if family == "deepseek_v2":
        return T.DeepseekV2ForCausalLM(
            T.DeepseekV2Config(**TINY, **MLA, n_routed_experts=8, num_experts_per_tok=3, moe_intermediate_size=24,
                               n_shared_experts=1, n_group=4, topk_group=2, first_k_dense_replace=1,
                               routed_scaling_factor=1.5))
    raise KeyError(family)


def _init(model):
    torch.manual_seed(0)
    for p in model.parameters():
        if p.dim() >= 2:
            torch.nn.init.normal_(p, std=0.05)
    for n, b in model.named_buffers():
        if "e_score_correction_bias" in n:
            b.normal_(std=0.1)
    return model.eval()


@pytest.mark.parametrize("family,n_blocks", [("mixtral", 2), ("qwen2_moe", 2), ("qwen3_moe", 2), ("deepseek_v3", 1),
                                             ("deepseek_v2", 1)])
def test_replaced_hf_model_matches_original(family, n_blocks):
    m = _init(_hf(family))
    ids = torch.randint(0, 100, (2, 9))
    with torch.no_grad():
        want = m(ids).logits
    m2 = copy.deepcopy(m)
    apply_autoep(m2, {"enabled": True, "autoep_size": 1})
    layers = [x for x in m2.modules() if isinstance(x, AutoEPMoELayer)]
    assert len(layers) == n_blocks and all(l._autoep_spec.model_family == {'qwen2_moe': 'qwen3_moe'}.get(family, family) for l in layers)
    with torch.no_grad():
        got = m2(ids).logits
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
    if family == "deepseek_v3":
        s = layers[0]._autoep_spec
        assert (s.score_func, s.num_expert_groups, s.num_limited_groups, s.route_scale) == ("sigmoid", 4, 2, 2.5)
        assert s.has_shared_experts and layers[0].router.e_score_correction_bias is not None
    if family == "qwen2_moe":
        assert layers[0].shared_experts_gate is not None and layers[0].router.route_norm is False


def test_router_logit_recorders_of_transformers_still_fire():
    """``output_router_logits=True`` (aux loss) works on the converted model: the original router module stays in place
    as ``router.gate`` and its forward hooks are fired with the tuple the HF router returns."""
    m = _init(_hf("mixtral"))
    ids = torch.randint(0, 100, (2, 9))
    with torch.no_grad():
        a = m(ids, output_router_logits=True, labels=ids)
    m2 = copy.deepcopy(m)
    apply_autoep(m2, P.AutoEPConfig(enabled=True, preset_model="mixtral"))
    with torch.no_grad():
        b = m2(ids, output_router_logits=True, labels=ids)
    assert len(b.router_logits) == 2 and abs(float(a.aux_loss) - float(b.aux_loss)) < 1e-6
    torch.testing.assert_close(b.router_logits[0], a.router_logits[0])


class _LegacyExpert(torch.nn.Module):

    def __init__(self, h, i):
        super().__init__()
        self.w1, self.w2, self.w3 = torch.nn.Linear(h, i, bias=False), torch.nn.Linear(i, h, bias=False), \
            torch.nn.Linear(h, i, bias=False)

    def forward(self, x):
        return self.w2(torch.nn.functional.silu(self.w1(x)) * self.w3(x))


class MixtralSparseMoeBlock(torch.nn.Module):
    """The pre-v5 Transformers block: ``ModuleList`` experts, returns ``(hidden, router_logits)``."""

    def __init__(self, h=16, i=24, e=4, k=2):
        super().__init__()
        self.gate = torch.nn.Linear(h, e, bias=False)
        self.experts = torch.nn.ModuleList([_LegacyExpert(h, i) for _ in range(e)])
        self.top_k = k

    def forward(self, x):
        b, s, h = x.shape
        flat = x.view(-1, h)
        logits = self.gate(flat)
        w, idx = torch.topk(torch.softmax(logits.float(), -1), self.top_k, -1)
        w = (w / w.sum(-1, keepdim=True)).to(x.dtype)
        out = torch.zeros_like(flat)
        for e, ex in enumerate(self.experts):
            t, slot = torch.where(idx == e)
            if t.numel():
                out.index_add_(0, t, ex(flat[t]) * w[t, slot, None])
        return out.view(b, s, h), logits


class _Wrap(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.block_sparse_moe = MixtralSparseMoeBlock()


def test_module_list_layout_and_custom_pattern():
    torch.manual_seed(0)
    m = _Wrap()
    x = torch.randn(2, 5, 16)
    want, want_logits = m.block_sparse_moe(x)
    cfg = P.AutoEPConfig(enabled=True, moe_layer_pattern=r"block_sparse_moe", expert_w1="w1", expert_w2="w2",
                         expert_w3="w3", top_k=2, return_router_logits=True)
    auto = AutoEP(m, cfg)
    specs = auto.ep_parser()
    assert len(specs) == 1 and specs[0].expert_storage == "module_list" and specs[0].ffn_hidden_size == 24
    auto.replace_moe_layers(specs)
    got, got_logits = m.block_sparse_moe(x)
    torch.testing.assert_close(got, want, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(got_logits, want_logits)
    with pytest.raises(ValueError, match="no MoE layers detected"):
        AutoEP(_Wrap(), P.AutoEPConfig(enabled=True, moe_layer_pattern=r"not_there")).ep_parser()


def test_router_group_limited_choice_bias_and_scaling():
    torch.manual_seed(0)
    r = TokenChoiceTopKRouter(8, 8, num_expert_groups=4, num_limited_groups=1, top_k=2, score_func="sigmoid",
                              route_norm=True, route_scale=3.0, group_score_func="max", dtype=torch.float32)
    x = torch.randn(32, 8)
    out = r(x)
    scores = torch.sigmoid(x @ r.weight.t())
    best_group = scores.view(32, 4, 2).max(-1).values.argmax(-1)
    assert torch.equal(out.selected_experts.sort(-1).values // 2, best_group[:, None].expand(-1, 2))  # one group only
    torch.testing.assert_close(out.top_scores.sum(-1), torch.full((32, ), 3.0))  # normalised, then scaled
    assert int(out.num_tokens_per_expert.sum()) == 64
    # a choice-only bias moves the selection, the weights still come from the unbiased scores
    plain = TokenChoiceTopKRouter(8, 8, top_k=1, score_func="softmax", route_norm=False, dtype=torch.float32)
    bias = torch.zeros(8)
    bias[5] = 10.0
    o = plain(x, expert_bias=bias)
    assert (o.selected_experts == 5).all()
    torch.testing.assert_close(o.top_scores[:, 0], torch.softmax(x @ plain.weight.t(), -1)[:, 5])
    for bad in (dict(num_expert_groups=3, num_limited_groups=1), dict(num_expert_groups=4, num_limited_groups=5),
                dict(num_expert_groups=4), dict(num_expert_groups=8, num_limited_groups=1)):
        with pytest.raises(ValueError):
            TokenChoiceTopKRouter(8, 8, top_k=2, **bad)
    # round-1 checkpoints stored the gate as ``router.weight``
    sd = {"weight": torch.ones(8, 8)}
    plain.load_state_dict(sd)
    assert float(plain.gate.weight.sum()) == 64.0


def test_score_apply_pre_scales_the_expert_input():
    torch.manual_seed(0)
    post = AutoEPMoELayer(16, 24, 4, 2, dtype=torch.float32)
    pre = AutoEPMoELayer(16, 24, 4, 2, dtype=torch.float32, score_apply="pre")
    pre.load_state_dict(post.state_dict())
    x = torch.randn(3, 7, 16)
    r = pre.router(x.view(-1, 16))
    want = torch.zeros(21, 16)
    for t in range(21):
        for w, e in zip(r.top_scores[t], r.selected_experts[t]):
            h = torch.nn.functional.linear(x.view(-1, 16)[t] * w, pre.experts.w13[e])
            want[t] += torch.nn.functional.linear(torch.nn.functional.silu(h[:24]) * h[24:], pre.experts.w2[e])
    torch.testing.assert_close(pre(x).view(21, 16), want, rtol=1e-4, atol=1e-6)
    assert not torch.allclose(pre(x), post(x))


def test_config_parsing_and_validation():
    c = P.parse_autoep_config({"enabled": True, "autoep_size": 2, "preset_model": "deepseek_v3", "expert_w3": None})
    assert c.autoep_size == 2 and c.expert_w3 is None and P.parse_autoep_config({}).expert_w3 is P._UNSET
    P.validate_autoep_config(c, world_size=4)
    with pytest.raises(ValueError, match="load_balance_coeff"):
        P.parse_autoep_config({"enabled": True, "load_balance_coeff": 0.001})
    for kw, msg in ((dict(preset_model="nope"), "Unknown preset_model"), (dict(score_apply="mid"), "score_apply"),
                    (dict(score_func="relu"), "score_func"), (dict(combine_impl="x"), "combine_impl"),
                    (dict(num_expert_groups=2, num_limited_groups=3), "num_limited_groups"),
                    (dict(expert_w1="a.b"), "no dots"), (dict(router_pattern=""), "non-empty"),
                    (dict(has_shared_experts="yes"), "boolean"), (dict(routed_scaling_factor="big"), "number"),
                    (dict(autoep_size=3), "must divide"), (dict(expert_tensor_parallel_size=2), "not supported")):
        with pytest.raises(ValueError, match=msg):
            P.validate_autoep_config(P.AutoEPConfig(enabled=True, **{"autoep_size": 2, **kw}), world_size=4)
    P.validate_autoep_config(P.AutoEPConfig(enabled=False, autoep_size=3), world_size=4)  # disabled: nothing checked
    # detection-time checks
    m = _init(_hf("mixtral"))
    with pytest.raises(ValueError, match="not divisible"):
        AutoEP(m, P.AutoEPConfig(enabled=True, autoep_size=3)).ep_parser()
    with pytest.raises(ValueError, match="exceeds num_experts"):
        AutoEP(m, P.AutoEPConfig(enabled=True, autoep_size=8)).ep_parser()
    with pytest.raises(ValueError, match="requires num_expert_groups"):
        AutoEP(m, P.AutoEPConfig(enabled=True, num_limited_groups=1)).ep_parser()


def test_preset_registry_and_adapters():
    assert set(P.available_preset_names()) >= {"mixtral", "qwen3_moe", "qwen3_5_moe", "deepseek_v2", "deepseek_v3"}
    assert P.preset_name_for_hf_model_type("qwen2_moe") == "qwen3_moe" and P.preset_name_for_hf_model_type("x") is None

    class C:
        model_type = "qwen3_5_moe"

    with pytest.raises(ValueError, match="text backbone"):
        P.resolve_preset_candidates(P.AutoEPConfig(enabled=True), C())
    C.model_type = "deepseek_v3"
    old = P.DeepSeekV3PresetAdapter.installed_transformers_version
    P.DeepSeekV3PresetAdapter.installed_transformers_version = lambda self: "4.40.0"
    try:
        with pytest.raises(ValueError, match="requires Transformers >= 5.0.0"):
            P.resolve_preset_candidates(P.AutoEPConfig(enabled=True), C())
    finally:
        P.DeepSeekV3PresetAdapter.installed_transformers_version = old
    base = P.get_preset("mixtral")
    over = P.apply_config_overrides(P.AutoEPConfig(expert_w1="up", expert_w3="gate2"), base)
    assert (over.expert_w1, over.expert_w3) == ("up", "gate2") and base.expert_w1 == "gate_up_proj"
    assert P.apply_config_overrides(P.AutoEPConfig(), base) is base
    v2 = P.get_preset_adapter("deepseek_v2")

    class V2:
        topk_method, n_group, topk_group, norm_topk_prob = "group_limited_greedy", 8, 3, True

    g = v2.resolve_group_routing(P.AutoEPConfig(), V2())
    assert (g.num_expert_groups, g.num_limited_groups, g.group_score_func) == (8, 3, "max")
    assert v2.resolve_route_norm(P.AutoEPConfig(), P.get_preset("deepseek_v2"), V2()) is False
    P.register_preset("mine", P.MoEModelPreset(**{**base.__dict__, "preset_adapter": "default"}))
    assert "mine" in P.available_preset_names()
    P.PRESET_MODELS.pop("mine")
    with pytest.raises(RuntimeError, match="unregistered adapter"):
        P.register_preset("bad", P.MoEModelPreset(**{**base.__dict__, "preset_adapter": "ghost"}))


def test_fused_expert_layout_classification():
    from deepspeed_b200.moe.fused_expert_layout import classify_fused_gate_up_layout as c
    a = c((4, 48, 16), (4, 16, 24))
    assert (a.layout, a.hidden_size, a.ffn_hidden_size, a.needs_transpose) == ("gate_up_first", 16, 24, False)
    b = c((4, 16, 48), (4, 24, 16))
    assert (b.layout, b.hidden_size, b.ffn_hidden_size, b.needs_transpose) == ("hidden_first", 16, 24, True)
    assert c((4, 16, 48), (4, 16, 16)) is None and c((16, 48), (4, 24, 16)) is None


def test_folding_spec_tables_and_policy():
    s = F_.build_folding_spec(world_size=8, tp_size=2, ep_size=4)
    assert (s.dp_size, s.edp_size, s.stage_size) == (4, 2, 8)
    t = F_.expected_folding_group_tables(s)
    assert t.tp_groups == ((0, 1), (2, 3), (4, 5), (6, 7))
    assert t.dense_dp_groups == ((0, 2, 4, 6), (1, 3, 5, 7))
    assert t.ep_groups == ((0, 1, 2, 3), (4, 5, 6, 7))
    assert t.edp_groups == ((0, 4), (1, 5), (2, 6), (3, 7))
    assert F_.local_folding_ranks(5, s) == {"tp": (4, 5), "dense_dp": (1, 3, 5, 7), "ep": (4, 5, 6, 7), "edp": (1, 5)}
    F_.assert_group_matches_spec(([[0, 1, 2, 3]], [[1, 5]]), s)
    with pytest.raises(RuntimeError):
        F_.assert_group_matches_spec(([[0, 2, 4, 6]], []), s)
    for kw in (dict(world_size=8, tp_size=3), dict(world_size=8, pp_size=3), dict(world_size=8, ep_size=3),
               dict(world_size=0), dict(world_size=8, tp_size=True)):
        with pytest.raises(ValueError):
            F_.build_folding_spec(**kw)
    F_.validate_folding_global(s, zero_stage=2)
    for kw in (dict(zero_stage=3), dict(sp_size=2), dict(deepcompile_enabled=True),
               dict(use_data_before_expert_parallel=True), dict(zero_offload_optimizer=True),
               dict(tp_preset="mixtral", ep_preset="qwen3_moe")):
        with pytest.raises(ValueError):
            F_.validate_folding_global(s, **kw)
    with pytest.raises(ValueError, match="autoep_size > 1"):
        F_.validate_folding_global(F_.build_folding_spec(world_size=8, tp_size=2, ep_size=1))
    with pytest.raises(ValueError, match="pp_size=1"):
        F_.validate_folding_global(F_.build_folding_spec(world_size=8, pp_size=2, tp_size=2, ep_size=2))


# ---- multi-rank ---------------------------------------------------------------------------------------------
def _hf_ep2(family):
    import deepspeed_b200.comm as dist
    m = _init(_hf(family)).train()
    torch.manual_seed(3)
    ids = torch.randint(0, 100, (2, 9))
    want = m(ids, labels=ids)
    want.loss.backward()
    rank = dist.get_rank()
    ref_block = [x for n, x in m.named_modules() if n.endswith(".mlp") and hasattr(x, "experts")][-1]
    ref_gate = ref_block.gate.weight.grad.clone()
    ref_up = ref_block.experts.gate_up_proj.grad.clone()
    m2 = _init(_hf(family)).train()
    apply_autoep(m2, {"enabled": True, "autoep_size": 2})
    got = m2(ids, labels=ids)  # the same batch on both ranks
    got.loss.backward()
    layer = [x for x in m2.modules() if isinstance(x, AutoEPMoELayer)][-1]
    nl = layer.num_local_experts
    assert layer.experts.w13.shape[0] == nl == ref_up.shape[0] // 2
    torch.testing.assert_close(got.loss, want.loss, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(layer.router.weight.grad, ref_gate, rtol=1e-4, atol=1e-6)
    # each local expert saw both ranks' (identical) tokens
    torch.testing.assert_close(layer.experts.w13.grad, 2 * ref_up[rank * nl:(rank + 1) * nl], rtol=1e-4, atol=1e-6)
    return True


@pytest.mark.parametrize("family", ["mixtral", "deepseek_v3"])
def test_hf_model_expert_parallel_two_ranks(family):
    assert all(run_distributed(_hf_ep2, 2, args=(family, )))


def _folded(rows):
    """tp=2 x ep=2 on two ranks: both lanes hold the same tokens, each routes half of them; outputs, input
    gradients and router gradients equal the unfolded single-process layer, expert gradients are the local slice
    (no factor)."""
    import deepspeed_b200.comm as dist
    from deepspeed_b200.utils import groups
    rank = dist.get_rank()
    torch.manual_seed(0)
    ref = AutoEPMoELayer(16, 24, 4, 2, dtype=torch.float32, aux_loss_coef=0.0)
    x = torch.randn(rows, 16)
    xr = x.clone().requires_grad_(True)
    yr = ref(xr)
    (yr * torch.arange(16.0)).sum().backward()
    spec = F_.build_folding_spec(world_size=2, tp_size=2, ep_size=2)
    F_.validate_folding_global(spec, zero_stage=1)
    h = F_.create_folding_groups(spec)
    assert h.tp_ranks == (0, 1) and h.ep_ranks == (0, 1) and h.edp_ranks == (rank, )
    groups._create_expert_and_data_parallel(2)
    lay = AutoEPMoELayer(16, 24, 4, 2, ep_size=2, dtype=torch.float32, tp_group=h.tp_group)
    sd = ref.state_dict()
    for k in ("experts.w13", "experts.w2"):
        sd[k] = sd[k][rank * 2:(rank + 1) * 2].clone()
    lay.load_state_dict(sd)
    lay.set_deepspeed_parallelism()
    xf = x.clone().requires_grad_(True)
    yf = lay(xf)
    (yf * torch.arange(16.0)).sum().backward()
    torch.testing.assert_close(yf, yr, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(xf.grad, xr.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lay.router.weight.grad, ref.router.weight.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(lay.experts.w13.grad, ref.experts.w13.grad[rank * 2:(rank + 1) * 2], rtol=1e-4,
                               atol=1e-6)
    torch.testing.assert_close(lay.experts.w2.grad, ref.experts.w2.grad[rank * 2:(rank + 1) * 2], rtol=1e-4, atol=1e-6)
    assert F_.autoep_folding_gradient_reduction_strategy(spec, lay.experts.w13) == "skip"
    F_.mark_autoep_folding_partial_router_parameter(lay.router.weight)
    assert F_.autoep_folding_gradient_reduction_strategy(spec, lay.router.weight, reduced_in_autograd=False) == "sum"
    return True


@pytest.mark.parametrize("rows", [12, 11])  # 11: the second lane carries one padding row
def test_autoep_folded_over_tensor_parallel(rows):
    assert all(run_distributed(_folded, 2, args=(rows, )))


def _engine_hf(zero_stage):
    """HF Mixtral through ``initialize`` with the ``expert_parallel`` section: blocks replaced, experts sharded over
    ep=2 and reduced over their own groups, training steps lower the loss identically on both ranks."""
    import deepspeed_b200 as ds
    m = _init(_hf("mixtral")).train()
    cfg = {"train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 1,
           "optimizer": {"type": "AdamW", "params": {"lr": 5e-3}},
           "zero_optimization": {"stage": zero_stage},
           "expert_parallel": {"enabled": True, "autoep_size": 2, "preset_model": "mixtral"}}
    engine, *_ = ds.initialize(model=m, config=cfg)
    layers = [x for x in engine.module.modules() if isinstance(x, AutoEPMoELayer)]
    assert len(layers) == 2 and engine.has_moe_layers and layers[0].experts.w13.shape[0] == 2
    assert layers[0].experts.w13.group_name == "ep_size_2" and layers[0].experts.w13.allreduce is False
    torch.manual_seed(11 + ds.comm.get_rank())
    ids = torch.randint(0, 100, (2, 9), device=engine.device)
    losses = []
    for _ in range(4):
        loss = engine(ids, labels=ids).loss
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0]
    return losses


@pytest.mark.parametrize("zero_stage", [0, 2])
def test_engine_applies_autoep_to_hf_model(zero_stage):
    out = run_distributed(_engine_hf, 2, args=(zero_stage, ))
    assert all(len(l) == 4 for l in out)


def _engine_rejects():
    import deepspeed_b200 as ds
    m = _init(_hf("mixtral"))
    with pytest.raises(ValueError, match="must divide"):
        ds.initialize(model=m, config={"train_micro_batch_size_per_gpu": 1,
                                       "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
                                       "expert_parallel": {"enabled": True, "autoep_size": 2}})
    return True


def test_engine_rejects_bad_expert_parallel_section():
    assert run_distributed(_engine_rejects, 1)[0]


def _autoep_checkpoint(tmp, zero_stage):
    """Expert-parallel checkpoints: every expert-parallel rank's experts and their optimizer state survive a save /
    load (per-expert files + per-rank optimizer state), training resumes bit-identically, and the per-expert files load
    at another expert-parallel degree."""
    import os
    import deepspeed_b200 as ds
    from deepspeed_b200.models.mixtral import MixtralConfig, MixtralForCausalLM

    def build(ep):
        torch.manual_seed(0)
        cfg = MixtralConfig.tiny(dtype=torch.float32, gradient_checkpointing=False, ep_size=ep,
                                 router_aux_loss_coef=0.0)
        dcfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
                "zero_optimization": {"stage": zero_stage}}
        return ds.initialize(model=MixtralForCausalLM(cfg), config=dcfg)[0]

    engine = build(2)
    rank = ds.comm.get_rank()
    torch.manual_seed(5 + rank)
    batches = [torch.randint(0, 512, (2, 16), device=engine.device) for _ in range(4)]

    def step(e, ids):
        loss = e(ids, labels=ids)
        e.backward(loss)
        e.step()
        return float(loss)

    for ids in batches[:2]:
        step(engine, ids)
    want = {k: v.clone() for k, v in engine.module.state_dict().items()}
    engine.save_checkpoint(tmp, tag="t")
    ref_losses = [step(engine, ids) for ids in batches[2:]]
    files = sorted(os.listdir(os.path.join(tmp, "t")))
    assert sum(f.startswith("layer_") and "_expert_" in f for f in files) == 2 * 4  # 2 MoE layers x 4 experts
    if zero_stage == 0:
        assert {"expp_rank_0_mp_rank_00_optim_states.pt", "expp_rank_1_mp_rank_00_optim_states.pt"} <= set(files)
    resumed = build(2)
    resumed.load_checkpoint(tmp, tag="t")
    for k, v in resumed.module.state_dict().items():
        assert torch.equal(v, want[k]), k
    got_losses = [step(resumed, ids) for ids in batches[2:]]
    assert got_losses == ref_losses, (got_losses, ref_losses)  # optimizer state came back too
    if zero_stage == 0:  # the per-expert files at another expert-parallel degree (weights only)
        single = build(1)
        single.load_checkpoint(tmp, tag="t", load_optimizer_states=False, load_module_only=True)
        sd = single.module.state_dict()
        k = "layers.0.block_sparse_moe.experts.w13"
        assert sd[k].shape[0] == 4 and torch.equal(sd[k][rank * 2:(rank + 1) * 2], want[k])
    return True


@pytest.mark.parametrize("zero_stage", [0, 1, 2])
def test_expert_parallel_checkpoint_roundtrip_and_resume(tmp_path, zero_stage):
    assert all(run_distributed(_autoep_checkpoint, 2, args=(str(tmp_path), zero_stage)))


def _autoep_zero_to_fp32(tmp):
    """zero_to_fp32 puts the experts of all expert-parallel ranks back together (they live in per-rank expert units, not
    in the data-parallel flats)."""
    import deepspeed_b200 as ds
    from deepspeed_b200.checkpoint.zero_to_fp32 import get_fp32_state_dict_from_zero_checkpoint
    from deepspeed_b200.models.mixtral import MixtralConfig, MixtralForCausalLM
    torch.manual_seed(0)
    cfg = MixtralConfig.tiny(dtype=torch.float32, gradient_checkpointing=False, ep_size=2, router_aux_loss_coef=0.0)
    engine = ds.initialize(model=MixtralForCausalLM(cfg),
                           config={"train_micro_batch_size_per_gpu": 2,
                                   "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
                                   "zero_optimization": {"stage": 2}})[0]
    rank = ds.comm.get_rank()
    torch.manual_seed(5 + rank)
    ids = torch.randint(0, 512, (2, 16), device=engine.device)
    loss = engine(ids, labels=ids)
    engine.backward(loss)
    engine.step()
    engine.save_checkpoint(tmp, tag="t")
    mine = {k: v.clone() for k, v in engine.module.state_dict().items()}
    ds.comm.barrier()
    full = get_fp32_state_dict_from_zero_checkpoint(tmp, tag="t")
    k = "layers.1.block_sparse_moe.experts.w2"
    assert full[k].shape[0] == 4 and torch.equal(full[k][rank * 2:(rank + 1) * 2], mine[k])
    dense = [n for n in mine if "experts" not in n and mine[n].dim() == 2][0]
    torch.testing.assert_close(full[dense], mine[dense])
    return True


def test_zero_to_fp32_reassembles_expert_parallel_weights(tmp_path):
    assert all(run_distributed(_autoep_zero_to_fp32, 2, args=(str(tmp_path), )))


def _autoep_universal(tmp, phase):
    """ep=2 ZeRO-1 checkpoint -> ds_to_universal -> resume at ep=1 (and ep=2): weights and Adam moments follow."""
    import os
    import deepspeed_b200 as ds
    from deepspeed_b200.models.mixtral import MixtralConfig, MixtralForCausalLM

    def build(ep, universal=False):
        torch.manual_seed(0)
        cfg = MixtralConfig.tiny(dtype=torch.float32, gradient_checkpointing=False, ep_size=ep,
                                 router_aux_loss_coef=0.0)
        dcfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
                "zero_optimization": {"stage": 1}, "checkpoint": {"load_universal": universal}}
        return ds.initialize(model=MixtralForCausalLM(cfg), config=dcfg)[0]

    rank = ds.comm.get_rank()
    if phase == "save":
        engine = build(2)
        torch.manual_seed(5 + rank)
        ids = torch.randint(0, 512, (2, 16), device=engine.device)
        for _ in range(2):
            loss = engine(ids, labels=ids)
            engine.backward(loss)
            engine.step()
        engine.save_checkpoint(tmp, tag="t")
        z = engine.optimizer
        moments = {}
        for (skey, a, b, gi, u, v) in z._segments:
            if u.group_key != "dp":
                moments[u.names[0]] = z.optimizer.state[v]["exp_avg"].clone()
        return {k: v.clone() for k, v in engine.module.state_dict().items() if "experts" in k}, moments
    engine = build(int(phase), universal=True)
    engine.load_checkpoint(tmp, tag="t_universal")
    sd = {k: v.clone() for k, v in engine.module.state_dict().items() if "experts" in k}
    z = engine.optimizer
    moments = {}
    for (skey, a, b, gi, u, v) in z._segments:
        if u.group_key != "dp":
            moments[u.names[0]] = z.optimizer.state[v]["exp_avg"].clone()
    return sd, moments


def test_universal_checkpoint_reshards_experts(tmp_path):
    from deepspeed_b200.checkpoint import ds_to_universal
    saved = run_distributed(_autoep_universal, 2, args=(str(tmp_path), "save"))
    args = ds_to_universal.parse_arguments(["--input_folder", str(tmp_path / "t"), "--output_folder",
                                            str(tmp_path / "t_universal"), "--num_extract_workers", "1",
                                            "--num_merge_workers", "1"])
    ds_to_universal.main(args)
    k = "layers.0.block_sparse_moe.experts.w13"
    want_full = torch.cat([saved[0][0][k], saved[1][0][k]], 0)
    one = run_distributed(_autoep_universal, 1, args=(str(tmp_path), "1"))[0]
    torch.testing.assert_close(one[0][k], want_full)
    two = run_distributed(_autoep_universal, 2, args=(str(tmp_path), "2"))
    for r in range(2):
        torch.testing.assert_close(two[r][0][k], saved[r][0][k])
        for name, m in saved[r][1].items():
            torch.testing.assert_close(two[r][1][name].view(-1), m.view(-1))


def test_dispatch_index_helpers():
    from deepspeed_b200.parallel.auto_ep import (combine_from_routed, count_tokens_per_expert, generate_permute_indices,
                                                 regroup_indices, resolve_combine_impl, resolve_score_apply_mode)
    torch.manual_seed(0)
    for _ in range(5):
        c = torch.randint(0, 5, (3, 4))
        seg, off = [], 0
        for r in range(3):
            for e in range(4):
                seg.append((e, r, off, int(c[r, e])))
                off += int(c[r, e])
        pieces = [torch.arange(o, o + n) for (e, r, o, n) in sorted(seg) if n > 0]
        want = torch.cat(pieces) if pieces else torch.empty(0, dtype=torch.long)
        assert torch.equal(regroup_indices(c), want)
    assert regroup_indices(torch.zeros(2, 2, dtype=torch.long)).numel() == 0
    c = torch.tensor([[2, 0, 3], [1, 4, 0]])
    idx, m_sizes, m_offsets = generate_permute_indices(c.reshape(-1), 3, 2, 16, 4)
    assert m_sizes.tolist() == [4, 4, 4] and m_offsets.tolist() == [4, 8, 12]
    assert idx.tolist() == [0, 1, 5, -1, 6, 7, 8, 9, 2, 3, 4, -1, -1, -1, -1, -1]
    assert count_tokens_per_expert(torch.tensor([[0, 2], [2, 3]]), 4).tolist() == [1, 0, 2, 1]
    out = combine_from_routed(torch.ones(4, 2), torch.tensor([0, 0, 1, 2]), torch.tensor([0.5, 0.25, 1.0, 2.0]), 3)
    assert out[:, 0].tolist() == [0.75, 1.0, 2.0]
    assert combine_from_routed(torch.ones(2, 2), torch.tensor([1, 1]), torch.ones(2), 2, "pre")[:, 0].tolist() == [0, 2]
    assert resolve_score_apply_mode("auto", "post") == "post" and resolve_score_apply_mode("pre", "post") == "pre"
    assert resolve_combine_impl("legacy_bmm") == "weighted_sum"
    with pytest.raises(ValueError):
        resolve_combine_impl("magic")


def test_repack_expert_weights_and_frozen_experts():
    from deepspeed_b200.parallel.auto_ep import repack_expert_requires_grad_flags, repack_expert_weights
    m = _init(_hf("mixtral"))
    for p in m.model.layers[0].mlp.experts.parameters():
        p.requires_grad_(False)
    auto = AutoEP(m, P.AutoEPConfig(enabled=True))
    specs = auto.ep_parser()
    src = m.model.layers[0].mlp.experts
    w1, w2, w3 = repack_expert_weights(src, specs[0], ep_rank=1, ep_size=2)
    assert w1.shape == (2, 48, 32) and w2.shape == (2, 32, 48) and w3.shape == (2, 48, 32)
    torch.testing.assert_close(torch.cat([w1, w3], 1), src.gate_up_proj[2:4])
    torch.testing.assert_close(w2, src.down_proj[2:4])
    assert repack_expert_requires_grad_flags(src, specs[0]) == (False, False)
    layers = auto.replace_moe_layers(specs)
    assert not layers[0].experts.w13.requires_grad and layers[1].experts.w13.requires_grad
    blk = _Wrap().block_sparse_moe
    cfg = P.AutoEPConfig(enabled=True, moe_layer_pattern=r"block_sparse_moe", expert_w1="w1", expert_w2="w2",
                         expert_w3="w3", top_k=2)
    w = _Wrap()
    spec = AutoEP(w, cfg).ep_parser()[0]
    a, b, c = repack_expert_weights(w.block_sparse_moe.experts, spec, 0, 2)
    assert a.shape == (2, 24, 16) and torch.equal(c[1], w.block_sparse_moe.experts[1].w3.weight)
    assert blk is not None
