"""Inference engines: blocked allocator, ragged engine == full-sequence forward, v1 engine wrapper
(reference tests/unit/inference/v2/ragged/*, test_inference.py)."""
import pytest
import torch


def test_blocked_allocator():
    from deepspeed_b200.inference.v2 import BlockedAllocator
    a = BlockedAllocator(8)
    x = a.allocate(3)
    assert a.free_blocks == 5 and len(set(x.tolist())) == 3
    a.free(x)
    assert a.free_blocks == 8
    with pytest.raises(ValueError):
        a.free(x[0].item())
    with pytest.raises(ValueError):
        a.allocate(9)


def test_ragged_engine_matches_full_forward():
    from deepspeed_b200.inference.v2 import InferenceEngineV2, RaggedInferenceEngineConfig
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    cfg = LlamaConfig.tiny(dtype=torch.float32, gradient_checkpointing=False)
    model = LlamaForCausalLM(cfg).eval()
    a, b = torch.randint(0, 512, (11, )), torch.randint(0, 512, (7, ))
    eng = InferenceEngineV2(model, RaggedInferenceEngineConfig(kv_block_size=4, num_kv_blocks=32))
    l1 = eng.put([1, 2], [a[:6], b[:7]])            # prefill (a is chunked)
    l2 = eng.put([1], [a[6:]])                      # second chunk of a
    with torch.no_grad():
        ref_a, ref_b = model(a[None]), model(b[None])
    torch.testing.assert_close(l1[1], ref_b[0, -1], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(l2[0], ref_a[0, -1], rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(l1[0], ref_a[0, 5], rtol=1e-4, atol=1e-4)
    free_before = eng.free_blocks
    eng.flush(1)
    eng.flush(2)
    assert eng.free_blocks > free_before and eng.free_blocks == 32


def test_init_inference_wrapper_and_generate():
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    model = LlamaForCausalLM(LlamaConfig.tiny(dtype=torch.float32, gradient_checkpointing=False))
    eng = ds.init_inference(model, dtype="fp32")
    ids = torch.randint(0, 512, (2, 5))
    out = eng.generate(ids, max_new_tokens=4)
    assert out.shape == (2, 9) and torch.equal(out[:, :5], ids)
    with torch.no_grad():
        nxt = model(ids)[:, -1].argmax(-1)
    assert torch.equal(out[:, 5], nxt)
