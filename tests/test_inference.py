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


def test_native_ragged_state_matches_python_bookkeeping():
    """C++ host runtime (csrc/cpu/ragged_host.cpp) vs the Python allocator / batch wrapper: same block tables, token
    descriptors and free-block accounting over a prefill + decode + flush scenario."""
    import random
    from deepspeed_b200.inference.v2.ragged import (BlockedKVCache, DSStateManager, NativeRaggedState,
                                                    RaggedBatchWrapper)
    bs, nb = 4, 40
    nat = NativeRaggedState(nb, bs, max_tokens=256, max_sequences=16)
    py = DSStateManager(BlockedKVCache(1, nb, bs, 1, 8, torch.float32, "cpu"))
    rng = random.Random(0)
    live = []
    for step in range(12):
        uids, lens = [], []
        for uid in list(live):
            if rng.random() < 0.7:
                uids.append(uid)
                lens.append(1 if rng.random() < 0.6 else rng.randint(2, 5))
        if step % 3 == 0:
            uids.append(100 + step)
            lens.append(rng.randint(5, 11))
            live.append(100 + step)
        if not uids:
            continue
        meta = nat.schedule(uids, lens, prefill_min_tokens=6, device=torch.device("cpu"))
        batch = RaggedBatchWrapper()
        for uid, n in zip(uids, lens):
            seq = py.get_or_create_sequence(uid)
            py.ensure_capacity(seq, n)
            seq.pre_forward(n)
            batch.insert_sequence(seq, torch.zeros(n, dtype=torch.long))
        tok_seq, tok_pos, table = batch.device_metadata(torch.device("cpu"))
        assert meta["tok_seq"].tolist() == tok_seq.tolist() and meta["tok_pos"].tolist() == tok_pos.tolist()
        # block tables agree up to the (equal) free-list order: compare block COUNTS and ownership disjointness
        for row, (seq, _, _, _) in enumerate(batch.seq_meta):
            got = [b for b in meta["table"][row].tolist()[:len(seq.blocks)]]
            assert len(got) == len(seq.blocks) == len(nat.blocks(seq.uid))
        for seq, _, _, _ in batch.seq_meta:
            seq.post_forward()
        assert nat.free_blocks == py.free_blocks
        n_p = sum(1 for n in lens if n >= 6)
        n_d = sum(1 for n in lens if n == 1)
        assert (meta["prefill"] is None) == (n_p == 0) and (meta["decode"] is None) == (n_d == 0)
        if meta["decode"] is not None:
            assert meta["decode"]["kv_len"].tolist() == [nat.seen_tokens(u) for u, n in zip(uids, lens) if n == 1]
        if step % 4 == 3 and live:
            gone = live.pop(0)
            nat.flush_sequence(gone)
            py.flush_sequence(gone)
            assert nat.free_blocks == py.free_blocks and nat.n_tracked_sequences == py.n_tracked_sequences
    all_blocks = [b for u in live for b in nat.blocks(u)]
    assert len(all_blocks) == len(set(all_blocks)) == nb - nat.free_blocks
    with pytest.raises(ValueError):
        nat.schedule([999], [bs * nb + 1], 6, torch.device("cpu"))
    assert nat.n_tracked_sequences == len(live)  # a refused batch changes nothing


def test_build_hf_engine_from_directory_and_serialized_checkpoint(tmp_path):
    """engine_factory: local HF directory -> engine; serialize() -> build_engine_from_ds_checkpoint round trip."""
    import os
    from transformers import LlamaConfig, LlamaForCausalLM
    from deepspeed_b200.inference import RaggedInferenceEngineConfig, build_hf_engine
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=128, hidden_size=64, intermediate_size=128, num_hidden_layers=2,
                      num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=64)
    hf = LlamaForCausalLM(cfg).eval()
    hf_dir, ds_dir = str(tmp_path / "hf"), str(tmp_path / "ds")
    hf.save_pretrained(hf_dir)
    ec = RaggedInferenceEngineConfig(state_manager={"max_ragged_batch_size": 64,
                                                    "memory_config": {"mode": "allocate", "size": 32}},
                                     tp={"tp_size": 1}, dtype=torch.float32)
    assert ec.num_kv_blocks == 32 and ec.state_manager.max_ragged_batch_size == 64 and ec.tensor_parallel.tp_size == 1
    eng = build_hf_engine(hf_dir, ec)
    ids = torch.randint(0, 128, (1, 9))
    with torch.no_grad():
        want = hf(ids).logits[0, -1]
    assert (eng.put([7], [ids[0]])[0].float().cpu() - want).abs().max() < 1e-4
    eng.serialize(ds_dir)
    assert sorted(os.listdir(ds_dir)) == ["ds_model_config.pkl", "metadata_rank_0_of_1.json", "params_rank_0_of_1.pt"]
    eng2 = build_hf_engine(ds_dir, RaggedInferenceEngineConfig(dtype=torch.float32))
    assert (eng2.put([1], [ids[0]])[0].float().cpu() - want).abs().max() < 1e-4
    with pytest.raises(FileNotFoundError):
        build_hf_engine(str(tmp_path / "missing"), ec)
