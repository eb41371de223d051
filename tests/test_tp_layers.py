"""Tensor-parallel layer zoo of ``parallel/layers.py`` (reference ``module_inject/layers.py``): named autograd
regions, gather / re-partition of layer parameters, Conv1D and Conv2d sharding, Yuan shared-QK mapping, norms."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from common import run_distributed
from deepspeed_b200.parallel import layers as L


@pytest.mark.parametrize("fn", [L.RowParallel, L.ColumnParallel, L.AsyncColumnParallel])
def test_named_regions_match_linear_on_one_rank(fn):
    torch.manual_seed(0)
    x = torch.randn(3, 5, 8, requires_grad=True)
    w = torch.randn(6, 8, requires_grad=True)
    b = torch.randn(6, requires_grad=True)
    y = fn.apply(None, x, w, b)
    y.square().sum().backward()
    got = [x.grad.clone(), w.grad.clone(), b.grad.clone()]
    x.grad = w.grad = b.grad = None
    yr = F.linear(x, w, b)
    yr.square().sum().backward()
    torch.testing.assert_close(y, yr)
    for a, r in zip(got, (x.grad, w.grad, b.grad)):
        torch.testing.assert_close(a, r, rtol=1e-5, atol=1e-5)


def test_rmsnorm_and_opt_embedding():
    torch.manual_seed(0)
    n = L.RMSNormalize(16, eps=1e-6)
    x = torch.randn(4, 16)
    torch.testing.assert_close(n(x), x / x.pow(2).mean(-1, keepdim=True).add(1e-6).sqrt())
    e = L.OPTEmbedding(weight=nn.Parameter(torch.arange(12.0).view(6, 2)))
    mask = torch.tensor([[0, 1, 1, 1], [1, 1, 1, 0]])
    out = e(mask)
    assert out.shape == (2, 4, 2)
    assert torch.equal(out[0, :, 0] / 2, torch.tensor([1.0, 2.0, 3.0, 4.0]))  # positions -1,0,1,2 offset by 2
    assert torch.equal(e(mask, past_key_values_length=2)[1, :, 0] / 2, torch.tensor([4.0, 1.0]))


def test_yuan_shared_qk_head_mapping():
    w = torch.arange(8.0).repeat_interleave(2).view(16, 1).expand(16, 4).contiguous()  # 8 heads of dim 2
    b = torch.arange(16.0)
    w0, b0 = L.shard_value_with_share_qk(w, b, 0, 2, True, num_kv_heads=8)
    w1, _ = L.shard_value_with_share_qk(w, b, 1, 2, True, num_kv_heads=8)
    assert w0[:, 0].tolist() == [0, 0, 1, 1, 4, 4, 5, 5] and w1[:, 0].tolist() == [2, 2, 3, 3, 6, 6, 7, 7]
    assert b0.tolist() == [0, 1, 2, 3, 8, 9, 10, 11]
    o, ob = L.shard_value_with_share_qk(w.t().contiguous(), torch.ones(4), 1, 2, False, num_kv_heads=8)
    assert o.shape == (4, 8) and ob.tolist() == [0.5] * 4
    with pytest.raises(RuntimeError):
        L.shard_value_with_share_qk(w, None, 0, 8, True, num_kv_heads=8)
    L.set_num_kv_heads(8)
    assert L.get_num_kv_heads() == 8 and L.shard_value_with_share_qk(w, None, 0, 2)[0].shape == (8, 4)
    L.set_num_kv_heads(None)


class _Conv1D(nn.Module):  # transformers.pytorch_utils.Conv1D

    def __init__(self, nf, nx):
        super().__init__()
        self.weight = nn.Parameter(torch.randn(nx, nf) * 0.1)
        self.bias = nn.Parameter(torch.randn(nf) * 0.1)

    def forward(self, x):
        return torch.addmm(self.bias, x.view(-1, x.size(-1)), self.weight).view(*x.shape[:-1], -1)


def _two_ranks():
    import deepspeed_b200.comm as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    group = dist.new_group([0, 1])
    torch.manual_seed(0)
    # Conv2d pair: output-channel shard then input-channel shard == the two full convolutions
    c1, c2 = nn.Conv2d(3, 7, 3, padding=1), nn.Conv2d(7, 5, 3, padding=1)
    x = torch.randn(2, 3, 8, 8)
    xr = x.clone().requires_grad_(True)
    want = c2(F.relu(c1(xr)))
    want.square().sum().backward()
    t1 = L.TensorParallelOcShardConv2d(c1, rank, world, group)
    t2 = L.TensorParallelIcShardConv2d(c2, rank, world, group)
    assert t1.conv.weight.shape[0] == (3 if rank == 0 else 4) and t2.conv.weight.shape[1] == t1.conv.weight.shape[0]
    xt = x.clone().requires_grad_(True)
    got = t2(F.relu(t1(xt)))
    got.square().sum().backward()
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(xt.grad, xr.grad, rtol=1e-4, atol=1e-5)
    lo, hi = t1.channel_range
    torch.testing.assert_close(t1.conv.weight.grad, c1.weight.grad[lo:hi], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(t2.conv.weight.grad, c2.weight.grad[:, lo:hi], rtol=1e-4, atol=1e-5)
    # uneven last-dim gather and its backward
    mine = torch.full((2, 3 if rank == 0 else 2), float(rank), requires_grad=True)
    full = L.GatherFromTensorParallelRegion.apply(mine, group, [3, 2])
    assert full.shape == (2, 5) and full[0].tolist() == [0, 0, 0, 1, 1]
    (full * torch.arange(5.0)).sum().backward()
    assert mine.grad[0].tolist() == ([0, 1, 2] if rank == 0 else [3, 4])
    # Conv1D (GPT-2) column + row pair, async column region, gather / re-partition of the layer's parameters
    a, b = _Conv1D(12, 8), _Conv1D(8, 12)
    h = torch.randn(4, 8)
    want2 = b(F.gelu(a(h)))
    col, row = L.conv_LinearLayer(a, group), L.Conv_LinearALlreduce(b, group)
    assert col.weight.shape == (6, 8) and row.weight.shape == (8, 6)
    torch.testing.assert_close(row(F.gelu(col(h))), want2, rtol=1e-4, atol=1e-5)
    hg = h.clone().requires_grad_(True)
    y = L.AsyncColumnParallel.apply(group, hg, col.weight, col.bias)
    y.sum().backward()
    full_w = a.weight.data.t()
    torch.testing.assert_close(hg.grad, full_w.sum(0).expand(4, 8), rtol=1e-4, atol=1e-5)  # summed over both shards
    shard = col.weight.data.clone()
    with L.GatherReplacedLayerParams(col, col):
        torch.testing.assert_close(col.weight.data, full_w)
        assert col.bias.shape == (12, )
    assert torch.equal(col.weight.data, shard) and col.bias.shape == (6, )
    # packed gate|up: both halves sharded separately
    lin = nn.Linear(4, 8, bias=False)
    lin.weight.data = torch.arange(8.0).view(8, 1).expand(8, 4).contiguous()
    gu = L.GateUpPack_LinearLayer(lin, group)
    assert gu.weight[:, 0].tolist() == ([0, 1, 4, 5] if rank == 0 else [2, 3, 6, 7])
    qkv = L.fused_LinearLayer(nn.Linear(4, 12, bias=False), group)
    assert qkv.weight.shape == (6, 4)
    wrapped = L.SubParamColumnParallel(L.FusedModuleWrapper(nn.Linear(4, 12), [8, 2, 2]), group)
    assert wrapped.weight.shape == (6, 4) and wrapped.bias.shape == (6, )
    return True


def test_tp_layers_two_ranks():
    assert all(run_distributed(_two_ranks, 2))


def test_tp_plan_converter_styles():
    from deepspeed_b200.parallel.autotp_config import PartitionType
    from deepspeed_b200.parallel.tp_plan_converter import TPPlanConverter
    specs = TPPlanConverter.convert({"layers.*.self_attn.q_proj": "colwise", "layers.*.mlp.down_proj": "rowwise",
                                     "lm_head": "colwise_gather_output", "layers.*.norm.weight":
                                     "replicated_with_grad_allreduce"})
    kinds = [(s.partition_type, s.gather_output, s.grad_allreduce) for s in specs]
    assert kinds == [(PartitionType.COLUMN, False, False), (PartitionType.ROW, False, False),
                     (PartitionType.COLUMN, True, False), (PartitionType.SKIP, False, True)]
    assert specs[0].patterns == [r".*layers\..*\.self_attn\.q_proj\.weight$"]
    assert specs[3].patterns == [r".*layers\..*\.norm\.weight$"]
    assert TPPlanConverter.convert({}) is None
    with pytest.raises(ValueError, match="unsupported partition style"):
        TPPlanConverter.convert({"a": "colwise", "b": "local_packed_rowwise"})


def _hf_plan():
    """HF Llama sharded by its own tp_plan through TpTrainingManager equals the unsharded model; lanes share RNG."""
    import transformers as T
    import deepspeed_b200.comm as dist
    from deepspeed_b200.parallel.tp import LinearAllreduce, LinearLayer
    from deepspeed_b200.runtime.tensor_parallel import TpTrainingManager
    torch.manual_seed(0)
    cfg = T.LlamaConfig(vocab_size=64, hidden_size=32, intermediate_size=48, num_hidden_layers=2,
                        num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=32)
    m = T.LlamaForCausalLM(cfg).eval()
    ids = torch.randint(0, 64, (2, 7))
    with torch.no_grad():
        want = m(ids).logits
    torch.manual_seed(100 + dist.get_rank())  # lanes start with different generators
    mgr = TpTrainingManager(m, tp_size=2, dtype=torch.float32)
    draws = torch.rand(4)
    both = [torch.empty(4) for _ in range(2)]
    dist.all_gather(both, draws)
    assert torch.equal(both[0], both[1])
    layer = m.model.layers[0]
    assert isinstance(layer.self_attn.q_proj, LinearLayer) and isinstance(layer.mlp.down_proj, LinearAllreduce)
    assert layer.self_attn.q_proj.weight.shape == (16, 32) and layer.mlp.down_proj.weight.shape == (32, 24)
    assert mgr.tp_config.tp_size == 2 and m.ds_autotp_parsed
    with torch.no_grad():
        got = m(ids).logits
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-5)
    return True


def test_hf_tp_plan_through_training_manager():
    assert all(run_distributed(_hf_plan, 2))


def test_fused_qkv_layouts_give_every_rank_whole_matching_heads():
    """Rows are labelled (kind, head); each layout's shard must hold, for every head it owns, all of that head's q,
    k and v rows (reference module_inject/fusedqkv_utils.py)."""
    from deepspeed_b200.parallel import fusedqkv_utils as Fq
    nh, hd, world = 8, 4, 2
    H = nh * hd

    def label(kinds_heads):  # list of (kind, head) per row block of hd rows
        return [(k, h) for k, h in kinds_heads for _ in range(hd)]

    def check(rows_by_rank, labels, replicated_kinds=()):
        seen = []
        for rows in rows_by_rank:
            got = [labels[i] for i in rows.tolist()]
            heads = {h for k, h in got if k == "q"}
            for kind in "kv":
                if kind in replicated_kinds:
                    continue
                assert {h for k, h in got if k == kind} == heads, (kind, heads)
            seen += [g for g in got if g[0] not in replicated_kinds]
        assert sorted(seen) == sorted(l for l in labels if l[0] not in replicated_kinds)  # a partition of the rows

    Fq.set_num_kv_heads(nh), Fq.set_num_attention_heads(nh), Fq.set_n_embd(H)
    bloom = label([(k, h) for h in range(nh) for k in "qkv"])
    check([Fq.fused_qkv_row_index("bloomtype", 3 * H, world, r) for r in range(world)], bloom)
    glm = label([(k, h) for k in "qkv" for h in range(nh)])
    check([Fq.fused_qkv_row_index("glmtype", 3 * H, world, r) for r in range(world)], glm)
    # codegen: 4 blocks of [q, v, k] sub-blocks, heads 2*blk, 2*blk+1 in block blk
    cg = label([(k, 2 * blk + i) for blk in range(4) for k in "qvk" for i in range(2)])
    rows = [Fq.fused_qkv_row_index("codegentype", 3 * H, world, r) for r in range(world)]
    check(rows, cg)
    assert all(len(r) == 3 * H // world for r in rows)
    # phi3: grouped KV heads (8 q heads, 2 kv heads): rank r owns q heads 4r..4r+3 and kv head r
    Fq.set_num_kv_heads(2)
    phi = label([("q", h) for h in range(nh)] + [("k", h) for h in range(2)] + [("v", h) for h in range(2)])
    for r in range(world):
        got = [phi[i] for i in Fq.fused_qkv_row_index("phi3type", H + 4 * hd, world, r, hidden_cols=H).tolist()]
        assert {h for k, h in got if k == "q"} == set(range(4 * r, 4 * r + 4))
        assert {h for k, h in got if k == "k"} == {r} == {h for k, h in got if k == "v"}
    # chatglm2: two KV groups behind n_embd q rows
    got = Fq.fused_qkv_row_index("glmtype", H + 4 * hd, world, 1)
    assert got.tolist() == list(range(16, 32)) + list(range(36, 40)) + list(range(44, 48))
    assert Fq.fused_qkv_subparam_sizes(torch.nn.Identity(), (H + 4 * hd, H)) == (H + 4 * hd, )
    # bigcode: q sharded, the single kv head replicated
    Fq.set_num_kv_heads(1)
    mq = label([("q", h) for h in range(nh)] + [("k", 0), ("v", 0)])
    rows = [Fq.fused_qkv_row_index("bigcodetype", H + 2 * hd, world, r) for r in range(world)]
    check(rows, mq, replicated_kinds=("k", "v"))
    assert all(r[-2 * hd:].tolist() == list(range(H, H + 2 * hd)) for r in rows)

    class FalconDecoderLayer(torch.nn.Module):

        def __init__(self):
            super().__init__()
            self.query_key_value = torch.nn.Linear(H, 3 * H)

    blk = FalconDecoderLayer()
    assert Fq.get_fused_qkv_type(blk) == "bloomtype" and Fq.get_fused_qkv_type(torch.nn.Linear(2, 2)) is None
    assert Fq.require_tp_fused_qkvw("h.0.self_attention.query_key_value", 2) and not Fq.require_tp_fused_qkvw("q", 2)
    Fq.set_num_kv_heads(nh)
    w = torch.arange(3 * H).float().view(-1, 1).expand(3 * H, 2)
    assert Fq.prepare_tp_fused_qkvw(blk, w, 2, 1)[:, 0].tolist() == list(range(48, 96))
    Fq.set_num_kv_heads(3)
    assert Fq.get_shard_size_list(96, 2, "self_attn.qkv") == [64, 32]  # 3 heads over 2 ranks: 2 + 1
    assert Fq.get_shard_size_list(96, 2, "mlp.up_proj") == [48, 48]
    parts = Fq.split_by_qkvlist_and_refuse([torch.arange(4), torch.arange(4) + 10, torch.arange(4) + 20], 2)
    assert parts[1].tolist() == [2, 3, 12, 13, 22, 23]
    for f, v in ((Fq.set_num_kv_heads, None), (Fq.set_num_attention_heads, None), (Fq.set_n_embd, None)):
        f(v)


def _hf_family_tp2(family):
    """A Transformers model sharded over two ranks by AutoTP reproduces the unsharded logits. Covers the fused-QKV
    layout families (interleaved heads, CodeGen blocks, multi-query, grouped KV), Conv1D projections, ALiBi models
    and tp_plan styles with gathered outputs."""
    import transformers as T
    from deepspeed_b200.config.sections import TPTrainingConfig
    from deepspeed_b200.parallel.tp import LinearAllreduce, LinearLayer, apply_autotp
    from deepspeed_b200.utils import groups
    torch.manual_seed(0)
    c = dict(vocab_size=64)
    small = dict(hidden_size=32, intermediate_size=64, num_hidden_layers=2, num_attention_heads=4,
                 num_key_value_heads=2, max_position_embeddings=32)
    build = {
        "gpt_neox": lambda: T.GPTNeoXForCausalLM(T.GPTNeoXConfig(**c, hidden_size=32, intermediate_size=64,
                                                                 num_hidden_layers=2, num_attention_heads=4,
                                                                 max_position_embeddings=32)),
        "bloom": lambda: T.BloomForCausalLM(T.BloomConfig(**c, hidden_size=32, n_layer=2, n_head=4)),
        "gpt_bigcode": lambda: T.GPTBigCodeForCausalLM(T.GPTBigCodeConfig(**c, n_embd=32, n_layer=2, n_head=4,
                                                                          n_positions=32, multi_query=True)),
        "gpt2": lambda: T.GPT2LMHeadModel(T.GPT2Config(**c, n_embd=32, n_layer=2, n_head=4, n_positions=32)),
        "opt": lambda: T.OPTForCausalLM(T.OPTConfig(**c, hidden_size=32, ffn_dim=64, num_hidden_layers=2,
                                                    num_attention_heads=4, max_position_embeddings=32,
                                                    word_embed_proj_dim=32)),
        "gptj": lambda: T.GPTJForCausalLM(T.GPTJConfig(**c, n_embd=32, n_layer=2, n_head=4, n_positions=32,
                                                       rotary_dim=4)),
        "falcon_gqa": lambda: T.FalconForCausalLM(T.FalconConfig(**c, hidden_size=32, num_hidden_layers=2,
                                                                 num_attention_heads=4, num_kv_heads=2,
                                                                 new_decoder_architecture=True)),
        "falcon_mqa": lambda: T.FalconForCausalLM(T.FalconConfig(**c, hidden_size=32, num_hidden_layers=2,
                                                                 num_attention_heads=4, multi_query=True,
                                                                 new_decoder_architecture=False, parallel_attn=True)),
        "phi3": lambda: T.Phi3ForCausalLM(T.Phi3Config(**c, **small, pad_token_id=0)),
        "qwen2": lambda: T.Qwen2ForCausalLM(T.Qwen2Config(**c, **small)),
        "gemma": lambda: T.GemmaForCausalLM(T.GemmaConfig(**c, **small, head_dim=8)),
        "mpt": lambda: T.MptForCausalLM(T.MptConfig(**c, d_model=32, n_heads=4, n_layers=2, max_seq_len=32)),
        "codegen": lambda: T.CodeGenForCausalLM(T.CodeGenConfig(**c, n_embd=64, n_layer=2, n_head=8, n_positions=32,
                                                                rotary_dim=4)),
    }
    m = build[family]().eval()
    ids = torch.randint(0, 64, (2, 7))
    with torch.no_grad():
        want = m(ids).logits
    groups.initialize(tp_size=2, force=True)
    apply_autotp(m, TPTrainingConfig(autotp_size=2, dtype=torch.float32), groups._get_model_parallel_group())
    n_tp = sum(isinstance(x, (LinearLayer, LinearAllreduce)) for x in m.modules())
    assert n_tp >= 8, n_tp  # 2 layers x (qkv / out / mlp in / mlp out) at least
    if family in ("gpt_neox", "bloom", "gpt_bigcode", "falcon_gqa", "falcon_mqa", "codegen"):
        qkv = [x for n, x in m.named_modules() if n.endswith(("query_key_value", "c_attn", "qkv_proj"))][0]
        assert qkv.row_index is not None
    with torch.no_grad():
        got = m(ids).logits
    torch.testing.assert_close(got, want, rtol=1e-4, atol=1e-4)
    return True


@pytest.mark.parametrize("family", ["gpt_neox", "bloom", "gpt_bigcode", "gpt2", "opt", "gptj", "falcon_gqa",
                                    "falcon_mqa", "phi3", "qwen2", "gemma", "mpt", "codegen"])
def test_autotp_hf_families_two_ranks(family):
    assert all(run_distributed(_hf_family_tp2, 2, args=(family, )))


def test_tp_init_utils(tmp_path):
    import base64
    import json
    import deepspeed_b200 as ds
    from deepspeed_b200.runtime.tensor_parallel import init_utils as U
    assert U.infer_config_dtype({"bf16": {"enabled": True}}) is torch.bfloat16
    assert U.infer_config_dtype({"fp16": {"enabled": True}}) is torch.float16 and U.infer_config_dtype({}) is None
    p = tmp_path / "c.json"
    p.write_text('{"train_batch_size": 4}')
    assert U.load_ds_config(str(p)) == {"train_batch_size": 4} and U.load_ds_config({"a": 1}) == {"a": 1}
    enc = base64.urlsafe_b64encode(json.dumps({"x": 2}).encode()).decode()
    assert U.load_ds_config(enc) == {"x": 2}
    with pytest.raises(ValueError):
        U.load_ds_config("definitely not a config")
    saved = ds._TP_MODEL_INIT_ARGS
    try:
        ds._TP_MODEL_INIT_ARGS = None
        U.record_tp_model_init_args(2, torch.bfloat16, None, None)
        U.record_tp_model_init_args(2, torch.bfloat16, None, None)
        with pytest.raises(ValueError, match="Conflicting"):
            U.record_tp_model_init_args(4, torch.bfloat16, None, None)
        with pytest.raises(ValueError, match="Conflicting"):
            U.record_tp_model_init_args(2, torch.bfloat16, object(), None)
    finally:
        ds._TP_MODEL_INIT_ARGS = saved
    assert U.tp_group_world_size(None, None) is None
