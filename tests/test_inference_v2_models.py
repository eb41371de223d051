"""FastGen model families: every policy is checked against the HuggingFace implementation of that family (tiny
random-init config, fp32 logits), then through the ragged engine (prefill + decode over the blocked KV cache).
Mirrors reference tests/unit/inference/v2/model_implementations."""
import pytest
import torch

T = pytest.importorskip("transformers")

from deepspeed_b200.inference.v2.model_implementations import REFERENCE_FAMILIES, SPECS, build_hf_model  # noqa: E402

TINY = dict(vocab_size=96, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
            max_position_embeddings=64)


def _cases():
    return {
        "llama": (T.LlamaConfig, T.LlamaForCausalLM, dict(num_key_value_heads=2)),
        "mistral": (T.MistralConfig, T.MistralForCausalLM, dict(num_key_value_heads=2, sliding_window=None)),
        "qwen2": (T.Qwen2Config, T.Qwen2ForCausalLM, dict(num_key_value_heads=2)),
        "phi3": (T.Phi3Config, T.Phi3ForCausalLM, dict(num_key_value_heads=2, pad_token_id=0)),
        "phi": (T.PhiConfig, T.PhiForCausalLM, dict()),
        "falcon": (T.FalconConfig, T.FalconForCausalLM, dict(new_decoder_architecture=False, multi_query=True,
                                                             parallel_attn=True, alibi=False, bias=False)),
        "falcon_new": (T.FalconConfig, T.FalconForCausalLM, dict(new_decoder_architecture=True, num_kv_heads=2,
                                                                 parallel_attn=True, alibi=False, bias=True)),
        "opt": (T.OPTConfig, T.OPTForCausalLM, dict(ffn_dim=64, word_embed_proj_dim=32)),
        "mixtral": (T.MixtralConfig, T.MixtralForCausalLM, dict(num_key_value_heads=2, num_local_experts=4)),
        "qwen2_moe": (T.Qwen2MoeConfig, T.Qwen2MoeForCausalLM,
                      dict(num_key_value_heads=2, num_experts=4, num_experts_per_tok=2, moe_intermediate_size=16,
                           shared_expert_intermediate_size=32)),
        "exaone4": (T.Exaone4Config, T.Exaone4ForCausalLM, dict(num_key_value_heads=2, sliding_window=None,
                                                                   layer_types=["full_attention"] * 2)),
    }


@pytest.mark.parametrize("name", list(_cases()))
def test_family_matches_huggingface(name):
    cfg_cls, model_cls, extra = _cases()[name]
    torch.manual_seed(0)
    hf = model_cls(cfg_cls(**{**TINY, **extra})).eval().float()
    with torch.no_grad():  # make every bias / norm weight non-trivial
        for n, p in hf.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    ours = build_hf_model(hf, dtype=torch.float32)
    ids = torch.randint(0, 96, (2, 11))
    with torch.no_grad():
        want = hf(input_ids=ids).logits
        got = ours(ids)
    torch.testing.assert_close(got, want, rtol=2e-4, atol=2e-4)


def test_every_reference_family_has_a_policy():
    assert set(REFERENCE_FAMILIES.values()) <= set(SPECS)
    assert len(REFERENCE_FAMILIES) >= 12


@pytest.mark.parametrize("name", ["llama", "phi", "opt", "mixtral"])
def test_ragged_engine_matches_dense(name):
    from deepspeed_b200.inference.v2.engine_v2 import InferenceEngineV2, RaggedInferenceEngineConfig
    cfg_cls, model_cls, extra = _cases()[name]
    torch.manual_seed(1)
    ours = build_hf_model(model_cls(cfg_cls(**{**TINY, **extra})).eval().float(), dtype=torch.float32)
    eng = InferenceEngineV2(ours, RaggedInferenceEngineConfig(kv_block_size=4, num_kv_blocks=64))
    a, b = torch.randint(0, 96, (9, )), torch.randint(0, 96, (5, ))
    with torch.no_grad():
        dense_a, dense_b = ours(a[None])[0], ours(b[None])[0]
    # prefill a in two chunks mixed with b's prefill, then one decode step each
    l1 = eng.put([0, 1], [a[:6], b[:4]])
    l2 = eng.put([0, 1], [a[6:], b[4:]])
    torch.testing.assert_close(l1[0], dense_a[5], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(l1[1], dense_b[3], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(l2[0], dense_a[8], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(l2[1], dense_b[4], rtol=1e-4, atol=1e-4)
    eng.flush(0)
    eng.flush(1)
    assert eng.free_blocks == 64
