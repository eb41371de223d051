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
