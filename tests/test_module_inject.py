"""``deepspeed.module_inject`` surface: BERT layer replacement / reversal, tensor slicing, declarative AutoTP specs.
Reference: module_inject/replace_module.py, autotp_config.py; tests/unit/model_parallelism/test_autotp_custom_patterns.py."""
import pytest
import torch

from common import run_distributed


def _bert(layers=2):
    from transformers import BertConfig, BertModel
    cfg = BertConfig(vocab_size=100, hidden_size=64, num_hidden_layers=layers, num_attention_heads=4,
                     intermediate_size=128, max_position_embeddings=32, hidden_dropout_prob=0.0,
                     attention_probs_dropout_prob=0.0)
    torch.manual_seed(0)
    return cfg, BertModel(cfg).eval()


def test_replace_and_revert_bert_layers_keep_the_function():
    import deepspeed_b200 as ds
    from transformers.models.bert.modeling_bert import BertLayer
    from deepspeed_b200.ops.transformer import DeepSpeedTransformerLayer
    cfg, model = _bert()
    ids = torch.randint(0, 100, (2, 16))
    with torch.no_grad():
        want = model(ids).last_hidden_state
    ds.replace_transformer_layer(BertLayer, model, model_config=cfg)
    assert sum(isinstance(m, DeepSpeedTransformerLayer) for m in model.modules()) == 2
    assert not any(isinstance(m, BertLayer) for m in model.modules())
    with torch.no_grad():
        got = model(ids).last_hidden_state
    assert torch.allclose(got, want, atol=2e-5), (got - want).abs().max()
    ds.revert_transformer_layer(BertLayer, model, cfg)
    assert sum(isinstance(m, BertLayer) for m in model.modules()) == 2
    with torch.no_grad():
        back = model(ids).last_hidden_state
    assert torch.allclose(back, want, atol=2e-5)


def test_tensor_slicing_and_group_quantizer():
    from deepspeed_b200.parallel import GroupQuantizer, ReplaceWithTensorSlicing
    sl = ReplaceWithTensorSlicing(mp_group=None, mp_size=2)
    sl.gpu_index = 1
    full = torch.arange(8 * 6, dtype=torch.float32).view(8, 6)
    out = sl.copy(torch.empty(4, 6), full)
    assert torch.equal(out, full[4:8])
    out = sl.copy(torch.empty(8, 3), full)
    assert torch.equal(out, full[:, 3:6])
    assert torch.equal(sl.copy(torch.empty(4), torch.arange(8.0)), torch.arange(4.0, 8.0))
    # fused q|k|v stacked on the last dim of a [in, 3*out] tensor: rank 1 takes the second half of each part
    qkv = torch.arange(4 * 12, dtype=torch.float32).view(4, 12)
    out = sl.strided_copy(torch.empty(4, 6), qkv, num_splits=3)
    assert torch.equal(out, torch.cat([qkv[:, 2:4], qkv[:, 6:8], qkv[:, 10:12]], dim=1))
    bias = torch.arange(12.0)
    out = sl.strided_copy(torch.empty(6), bias, num_splits=3)
    assert torch.equal(out, torch.tensor([2., 3., 6., 7., 10., 11.]))
    w = torch.randn(16, 32)
    q = GroupQuantizer(q_int8=True, group_size=4).quantize(w)
    assert q.dtype == torch.int8 and q.scale.shape[0] == 4
    group_scale = q.scale.reshape(3, 4)[0]  # reference layout: [whole tensor | first half | second half] x groups
    deq = q.float().reshape(4, -1) * group_scale[:, None]
    assert (deq.reshape(16, 32) - w).abs().max() < w.abs().max() / 100


def test_tp_layer_spec_validation_and_matching():
    from deepspeed_b200.parallel import AutoTPConfig, AutoTPPresets, PartitionType, TPLayerSpec, merge_autotp_configs
    s = TPLayerSpec(patterns=[r".*\.qkv_proj\.weight$"], shape=((8, 2, 2), -1), partition_dim=0)
    assert s.has_unequal_sub_params() and s.get_sub_param_sizes() == (8, 2, 2) and s.get_num_sub_params() == 3
    assert s.sub_param_sizes_for(12) == [8, 2, 2]
    with pytest.raises(ValueError):
        s.sub_param_sizes_for(13)
    e = TPLayerSpec(patterns=["gate_up"], shape=(2, -1))
    assert e.sub_param_sizes_for(10) == [5, 5] and e.get_partition_dim() == 0
    assert TPLayerSpec(patterns=["x"], partition_type="row").get_partition_dim() == 1
    with pytest.raises(ValueError):
        TPLayerSpec(patterns=["x"], shape=((1, 2), (3, 4)))
    with pytest.raises(ValueError):
        TPLayerSpec(patterns=["x"], shape=(-1, (3, 4)), partition_dim=0)
    only_llama = TPLayerSpec(patterns=["proj"], model_types=["llama"])
    assert only_llama.matches("a.proj.weight", "llama") and not only_llama.matches("a.proj.weight", "bloom")
    cfg = AutoTPConfig.from_dict({"use_default_specs": False, "layer_specs": [
        {"patterns": r".*\.router\.weight$", "partition_type": "skip"},
        {"patterns": [r".*\.down\.weight$"], "partition_type": "row"}]})
    assert cfg.find_matching_spec("m.router.weight").partition_type == PartitionType.SKIP
    assert cfg.find_matching_spec("m.other.weight") is None and not cfg.use_default_specs
    llama = AutoTPPresets.get_preset("llama")
    assert llama.find_matching_spec("model.layers.0.self_attn.o_proj.weight").partition_type == PartitionType.ROW
    assert AutoTPPresets.get_preset("no-such-model") is None
    merged = merge_autotp_configs(llama, cfg)
    assert merged.layer_specs[0] is cfg.layer_specs[0] and len(merged.layer_specs) == 2 + len(llama.layer_specs)


class _Toy(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.up = torch.nn.Linear(16, 32, bias=False)       # fused [a | b] halves, column parallel
        self.router = torch.nn.Linear(16, 4, bias=False)    # must stay replicated
        self.down = torch.nn.Linear(16, 16, bias=False)     # row parallel
        self.odd = torch.nn.Linear(16, 16, bias=False)      # matches nothing

    def forward(self, x):
        a, b = self.up(x).chunk(2, dim=-1)
        return self.down(a * torch.sigmoid(b)) + self.odd(x) + self.router(x).sum(-1, keepdim=True)


def _custom_specs(strict):
    import deepspeed_b200 as ds
    from deepspeed_b200.config.sections import TPTrainingConfig
    from deepspeed_b200.parallel import LinearAllreduce, LinearLayer, apply_autotp
    from deepspeed_b200.utils import groups
    torch.manual_seed(0)
    ref = _Toy()
    x = torch.randn(3, 16)
    want = ref(x)
    model = _Toy()
    model.load_state_dict(ref.state_dict())
    ds.init_distributed()
    groups.initialize(tp_size=2, force=True)
    pc = {"use_default_specs": False, "strict_mode": strict, "layer_specs": [
        {"patterns": [r"^up\.weight$"], "partition_type": "column", "shape": [2, -1], "partition_dim": 0},
        {"patterns": [r"^router\.weight$"], "partition_type": "skip"},
        {"patterns": [r"^down\.weight$"], "partition_type": "row"}]}
    try:
        apply_autotp(model, TPTrainingConfig(autotp_size=2, partition_config=pc), groups._get_model_parallel_group())
    except ValueError as e:
        return "strict", str(e)
    kinds = (type(model.up).__name__, type(model.router).__name__, type(model.down).__name__, type(model.odd).__name__)
    got = model(x)
    return kinds, float((got - want).abs().max()), tuple(model.up.weight.shape)


def test_autotp_follows_layer_specs():
    kinds, err, up_shape = run_distributed(_custom_specs, 2, args=(False, ))[0]
    assert kinds == ("LinearLayer", "Linear", "LinearAllreduce", "Linear")
    assert up_shape == (16, 16) and err < 1e-5
    tag, msg = run_distributed(_custom_specs, 2, args=(True, ))[0]
    assert tag == "strict" and "odd" in msg


def test_merge_tp_model_init_conflicts():
    import deepspeed_b200 as ds
    ds._TP_MODEL_INIT_ARGS = dict(tp_size=2, dtype=torch.bfloat16, config=None)
    try:
        cfg = {}
        ds.merge_tp_model_init_into_config(cfg, None, None, None)
        assert cfg["tensor_parallel"]["autotp_size"] == 2 and cfg["tensor_parallel"]["tp"]["tp_size"] == 2
        with pytest.raises(ValueError):
            ds.merge_tp_model_init_into_config({"tensor_parallel": {"autotp_size": 4}}, None, None, None)
        with pytest.raises(ValueError):
            ds.merge_tp_model_init_into_config({"tensor_parallel": {"tp": {"tp_size": 8}}}, None, None, None)
    finally:
        ds._TP_MODEL_INIT_ARGS = None


def test_top_level_reference_names():
    import deepspeed_b200 as ds
    for n in """ADAM_OPTIMIZER LAMB_OPTIMIZER MUON_OPTIMIZER DeepSpeedDataLoader DeepSpeedInferenceConfig
    DeepSpeedOptimizer OnDevice TORCH_DISTRIBUTED_DEFAULT_PORT domino is_compile_supported git_hash git_branch
    replace_transformer_layer revert_transformer_layer set_autotp_mode merge_tp_model_init_into_config
    DeepSpeedEngine PipelineEngine InferenceEngine DeepSpeedHybridEngine PipelineModule checkpointing zero
    DeepSpeedTransformerLayer DeepSpeedTransformerConfig init_distributed initialize init_inference tp_model_init
    add_config_arguments""".split():
        assert hasattr(ds, n), n
    from deepspeed_b200.runtime.base_optimizer import ZeROOptimizer
    from deepspeed_b200.runtime.zero import ZeroShardedOptimizer
    assert issubclass(ZeroShardedOptimizer, ZeROOptimizer) and issubclass(ZeROOptimizer, ds.DeepSpeedOptimizer)
    assert ds.is_compile_supported()


def test_generic_class_keyed_replacement(tmp_path):
    import torch
    import torch.nn as nn
    from deepspeed_b200.parallel.replace_module import (get_transformer_name, quantize_module, replace_module,
                                                        skip_level_0_prefix)

    class Block(nn.Module):

        def __init__(self):
            super().__init__()
            self.fc = nn.Linear(4, 4)

    class GPTModel(nn.Module):

        def __init__(self):
            super().__init__()
            self.transformer = nn.Module()
            self.transformer.h = nn.ModuleList([Block(), Block()])
            self.head = nn.Linear(4, 4)

    m = GPTModel()
    seen = []

    def fn(child, policy, layer_id):
        seen.append((policy, layer_id))
        return nn.Identity()

    replace_module(m, Block, fn, "mypolicy")
    assert seen == [("mypolicy", 0), ("mypolicy", 1)] and isinstance(m.transformer.h[1], nn.Identity)
    torch.save({"k": torch.ones(1)}, tmp_path / "ck.pt")
    got = {}

    def fn2(child, policy, layer_id, prefix="", state_dict=None):
        got[prefix] = state_dict["k"].item()
        return child

    replace_module(m, nn.Linear, fn2, None, checkpoint=str(tmp_path / "ck.pt"))
    assert got == {"head": 1.0}
    quantize_module(m, nn.Linear, lambda lin: nn.ReLU())
    assert isinstance(m.head, nn.ReLU)
    assert get_transformer_name(GPTModel()) == "transformer.h"

    class BloomModel(nn.Module):
        pass

    assert skip_level_0_prefix(BloomModel(), {"h.0.w": 1}) and not skip_level_0_prefix(BloomModel(), {"model.h": 1})
    assert not skip_level_0_prefix(GPTModel(), None)
