"""Domino TP layer (two micro-batches, async all-reduces) against the same layer computed without tensor parallelism.
Mirrors reference tests/unit/runtime/domino (forward/backward parity)."""
from types import SimpleNamespace

import torch

from common import run_distributed


def _domino():
    import torch.distributed as dist
    from deepspeed_b200.runtime.domino import DominoTransformerLayer
    rank, world = dist.get_rank(), dist.get_world_size()
    cfg = SimpleNamespace(hidden_size=32, num_attention_heads=4, ffn_hidden_size=64, num_layers=1)
    torch.manual_seed(0)
    full = DominoTransformerLayer(cfg, tp_group=None)            # world-1 reference (tp_world == 1)
    torch.manual_seed(0)
    tp = DominoTransformerLayer(cfg, tp_group=dist.group.WORLD)
    h, hd, nh = 32, 8, 4
    with torch.no_grad():  # shard the reference weights: heads / ffn columns split over ranks
        qkv_w = full.self_attention.query_key_value.weight.view(nh, 3 * hd, h)
        qkv_b = full.self_attention.query_key_value.bias.view(nh, 3 * hd)
        hs = slice(rank * nh // world, (rank + 1) * nh // world)
        tp.self_attention.query_key_value.weight.copy_(qkv_w[hs].reshape(-1, h))
        tp.self_attention.query_key_value.bias.copy_(qkv_b[hs].reshape(-1))
        cols = slice(rank * h // world, (rank + 1) * h // world)
        tp.self_attention.dense.weight.copy_(full.self_attention.dense.weight[:, cols])
        tp.self_attention.dense.bias.copy_(full.self_attention.dense.bias)
        f = slice(rank * 64 // world, (rank + 1) * 64 // world)
        tp.dense_h_to_4h.weight.copy_(full.dense_h_to_4h.weight[f])
        tp.dense_h_to_4h.bias.copy_(full.dense_h_to_4h.bias[f])
        tp.dense_4h_to_h.weight.copy_(full.dense_4h_to_h.weight[:, f])
        tp.dense_4h_to_h.bias.copy_(full.dense_4h_to_h.bias)
        for a, b in ((tp.input_layernorm, full.input_layernorm),
                     (tp.post_attention_layernorm, full.post_attention_layernorm)):
            a.load_state_dict(b.state_dict())
    torch.manual_seed(1)
    x = torch.randn(4, 6, 32)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya, yb = full(xa), tp(xb)
    ya.sum().backward()
    yb.sum().backward()
    return float((ya - yb).abs().max()), float((xa.grad - xb.grad).abs().max())


def test_domino_layer_matches_unsharded():
    for fwd, bwd in run_distributed(_domino, 2):
        assert fwd < 1e-5 and bwd < 1e-5, (fwd, bwd)
