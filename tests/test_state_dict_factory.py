"""Megatron checkpoint merge / split across inference TP degrees and int8 weight quantization
(reference tests/unit/checkpoint + runtime/state_dict_factory.py semantics)."""
import torch

from deepspeed_b200.runtime.state_dict_factory import SDLoaderFactory
from deepspeed_b200.runtime.weight_quantizer import WeightQuantization


def _full_sd(h=16, ffn=64):
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    return {
        "word_embeddings.weight": r(32, h),
        "layers.0.attention.query_key_value.weight": r(3 * h, h),
        "layers.0.attention.query_key_value.bias": r(3 * h),
        "layers.0.attention.dense.weight": r(h, h),
        "layers.0.mlp.dense_h_to_4h.weight": r(ffn, h),
        "layers.0.mlp.dense_h_to_4h.bias": r(ffn),
        "layers.0.mlp.dense_4h_to_h.weight": r(h, ffn),
        "layers.0.input_layernorm.weight": r(h),
    }


def _shard(full, mp, rank, version):
    out = {}
    for k, v in full.items():
        if "query_key_value" in k:
            if version == 0:  # [q | k | v] blocks, each split over ranks
                q, kk, vv = v.chunk(3, dim=0)
                out[k] = torch.cat([t.chunk(mp, dim=0)[rank] for t in (q, kk, vv)], dim=0)
            else:
                out[k] = v.chunk(mp, dim=0)[rank]
        elif "attention.dense.weight" in k or "dense_4h_to_h.weight" in k:
            out[k] = v.chunk(mp, dim=1)[rank]
        elif "dense_h_to_4h" in k or "word_embeddings" in k:
            out[k] = v.chunk(mp, dim=0)[rank]
        else:
            out[k] = v
    return out


def test_merge_and_split_roundtrip(tmp_path):
    full = _full_sd()
    for version in (0, 2.0):
        paths = []
        for r in range(4):
            p = str(tmp_path / f"v{version}_mp{r}.pt")
            torch.save({"module": _shard(full, 4, r, version), "mp_world_size": 4, "checkpoint_version": version}, p)
            paths.append(p)
        loader = SDLoaderFactory.get_sd_loader(paths, sd_type="Megatron", version=version)
        # 4 shards -> 2 ranks (merge)
        for rank in range(2):
            _, sd, (scales, merge_count) = loader.load(2, rank)
            want = _shard(full, 2, rank, version)
            assert merge_count == 2 and scales is None
            for k in want:
                torch.testing.assert_close(sd["module"][k], want[k])
        # 4 shards -> 8 ranks (split)
        for rank in (0, 5):
            _, sd, _ = loader.load(8, rank)
            want = _shard(full, 8, rank, version)
            for k in want:
                torch.testing.assert_close(sd["module"][k], want[k])
        # same degree: passthrough
        _, sd, _ = loader.load(4, 3)
        torch.testing.assert_close(sd["module"]["layers.0.attention.dense.weight"],
                                   _shard(full, 4, 3, version)["layers.0.attention.dense.weight"])


def test_weight_quantization_int8():
    torch.manual_seed(0)
    w = torch.randn(64, 16)
    q = WeightQuantization(mlp_extra_grouping=False, mp_size=1)
    (qi, ), = [q.Quantize([w.clone()], 8, 4, key="layers.0.mlp.dense_h_to_4h.weight")]
    assert qi.dtype == torch.int8
    scale = q.mlph4h_scales[0].view(-1)            # 1 / dequant-scale per group
    deq = (qi.float().flatten().view(4, -1) / scale.view(4, 1)).view_as(w)
    assert (deq - w).abs().max() < w.abs().max() / 100
    sd = {"layers.0.attention.query_key_value.weight": torch.randn(48, 16),
          "layers.0.attention.dense.weight": torch.randn(16, 16),
          "layers.0.mlp.dense_h_to_4h.weight": torch.randn(64, 16),
          "layers.0.mlp.dense_4h_to_h.weight": torch.randn(16, 64), "other": torch.randn(3)}
    out, scales = WeightQuantization(mp_size=1).sd_quantize_megatron(dict(sd), 8, 2)
    assert all(out[k].dtype == torch.int8 for k in sd if k != "other") and out["other"].dtype == torch.float32
    assert scales.shape[0] == 1 and scales.shape[1] == 4
