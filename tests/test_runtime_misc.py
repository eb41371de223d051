"""Small runtime features: progressive layer drop, eigenvalue, sparse tensors, MoQ quantizer, TiledLinear, utils.
Mirrors reference tests/unit/runtime/{test_pld.py,sparse_tensor/,zero/test_zero_tiled.py,utils/}."""
import math

import torch

from deepspeed_b200.runtime.eigenvalue import Eigenvalue
from deepspeed_b200.runtime.progressive_layer_drop import ProgressiveLayerDrop
from deepspeed_b200.runtime.quantize import Quantizer
from deepspeed_b200.runtime.sparse_tensor import SparseTensor
from deepspeed_b200.runtime.utils import (CheckOverflow, clip_grad_norm_, get_global_norm_of_tensors,
                                          partition_balanced, partition_uniform)
from deepspeed_b200.runtime.zero.tiling import TiledLinear


def test_pld_schedule():
    pld = ProgressiveLayerDrop(theta=0.5, gamma=0.01)
    assert pld.get_theta() == 1.0
    for step in (0, 10, 1000):
        pld.update_state(step)
        assert abs(pld.get_theta() - ((1 - 0.5) * math.exp(-0.01 * step) + 0.5)) < 1e-9
    assert pld.get_state() == {"progressive_layer_drop": True, "pld_theta": pld.get_theta()}


def test_sparse_tensor_roundtrip():
    dense = torch.zeros(10, 4)
    dense[2] = torch.arange(4.0)
    dense[7] = 1.5
    st = SparseTensor(dense)
    assert st.indices.tolist() == [2, 7]
    torch.testing.assert_close(st.to_dense(), dense)
    st.add(SparseTensor(dense))
    torch.testing.assert_close(st.to_dense(), 2 * dense)
    emb = torch.nn.Embedding(10, 4, sparse=True)
    emb(torch.tensor([1, 1, 3])).sum().backward()
    st2 = SparseTensor(emb.weight.grad)
    torch.testing.assert_close(st2.to_dense(), emb.weight.grad.to_dense())
    torch.testing.assert_close(st2.to_coo_tensor().to_dense(), emb.weight.grad.to_dense())


def test_eigenvalue_power_iteration():
    torch.manual_seed(0)

    class Block(torch.nn.Module):

        def __init__(self, scale):
            super().__init__()
            self.w = torch.nn.Parameter(torch.randn(6))
            self.scale = scale

        def forward(self, x):
            return x + self.scale * (self.w**2).sum()

    class Net(torch.nn.Module):

        def __init__(self):
            super().__init__()
            self.layers = torch.nn.ModuleList([Block(1.0), Block(4.0)])

        def forward(self, x):
            for layer in self.layers:
                x = layer(x)
            return x

    net = Net()
    net(torch.zeros(())).backward(create_graph=True)
    ev = Eigenvalue(max_iter=50, tol=1e-4, layer_name="layers", layer_num=2)
    out = ev.compute_eigenvalue(net)
    # Hessians are 2*scale*I -> top eigenvalues 2 and 8 -> normalised 0.25 and 1.0
    e0, e1 = out[id(net.layers[0].w)], out[id(net.layers[1].w)]
    assert e0[1] == 0 and e1[1] == 1
    assert abs(e0[0] - 0.25) < 1e-3 and abs(e1[0] - 1.0) < 1e-6


def test_moq_quantizer_schedule():
    torch.manual_seed(0)
    p = torch.nn.Parameter(torch.randn(8, 16))
    p.start_bits, p.target_bits, p.q_period = 8, 4, 2
    q = Quantizer(q_groups=2, q_type=0, q_rounding=0)
    orig = p.data.clone()
    for _ in range(40):
        q.quantize([[p]], overflow=False, eigenvalue_enabled=False)
    assert p.start_bits == 4
    levels = torch.unique(p.data.reshape(2, -1)[0]).numel()
    assert levels <= 16
    assert (p.data - orig).abs().max() < orig.abs().max()
    asym = Quantizer.fake_quantize(orig, 8, 4, symmetric=False)
    assert (asym - orig).abs().max() <= (orig.max() - orig.min()) / 255
    sto = torch.stack([Quantizer.fake_quantize(orig, 4, 1, stochastic=True) for _ in range(200)]).mean(0)
    assert (sto - orig).abs().mean() < (Quantizer.fake_quantize(orig, 4, 1) - orig).abs().mean()


def test_tiled_linear_matches_dense():
    torch.manual_seed(0)
    dense = torch.nn.Linear(24, 18)
    tiled = TiledLinear(24, 18, in_splits=3, out_splits=2, init_linear=dense)
    x = torch.randn(5, 24)
    torch.testing.assert_close(tiled(x), dense(x), rtol=1e-5, atol=1e-5)
    dense(x).sum().backward()
    tiled(x).sum().backward()
    torch.testing.assert_close(tiled.linears[1][2].weight.grad, dense.weight.grad[9:18, 16:24])
    parts = TiledLinear(24, 18, out_splits=3, combine_out_splits=False)(x)
    assert [p.shape[-1] for p in parts] == [6, 6, 6]


def test_norm_clip_partition_utils():
    ps = [torch.nn.Parameter(torch.ones(4)), torch.nn.Parameter(torch.ones(9))]
    for p in ps:
        p.grad = torch.full_like(p, 2.0)
    assert abs(get_global_norm_of_tensors([p.grad for p in ps]) - math.sqrt(13 * 4)) < 1e-5
    norm = clip_grad_norm_(ps, max_norm=1.0)
    assert abs(norm - math.sqrt(52)) < 1e-5
    assert abs(get_global_norm_of_tensors([p.grad for p in ps]) - 1.0) < 1e-4
    assert not CheckOverflow([ps]).check()
    ps[0].grad[1] = float("inf")
    assert CheckOverflow([ps]).check()
    assert partition_uniform(10, 3) == [0, 4, 7, 10] or partition_uniform(10, 3)[-1] == 10
    parts = partition_balanced([1, 1, 1, 10, 1, 1], 2)
    assert parts[0] == 0 and parts[-1] == 6 and len(parts) == 3
