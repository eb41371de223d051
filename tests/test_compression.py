"""Compression library end to end: init_compression -> scheduler -> training signal on mask scores -> redundancy_clean
preserves the function while shrinking dimensions. Mirrors reference tests/unit/compression/test_compression.py."""
import torch

from deepspeed_b200.compression import compression_scheduler, init_compression, redundancy_clean
from deepspeed_b200.compression.basic_layer import (Conv2dLayer_Compress, Embedding_Compress, LinearLayer_Compress,
                                                    BNLayer_Compress)
from deepspeed_b200.compression.config import get_compression_config
from deepspeed_b200.compression.utils import (AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer,
                                              TopKBinarizer)


class Attn(torch.nn.Module):

    def __init__(self, h=32):
        super().__init__()
        self.q, self.k, self.v, self.o = (torch.nn.Linear(h, h) for _ in range(4))

    def forward(self, x):
        return self.o(torch.tanh(self.q(x)) * self.k(x) + self.v(x))


class Block(torch.nn.Module):

    def __init__(self, h=32):
        super().__init__()
        self.attn, self.fc1, self.fc2 = Attn(h), torch.nn.Linear(h, 4 * h), torch.nn.Linear(4 * h, h)

    def forward(self, x):
        x = x + self.attn(x)
        return x + self.fc2(torch.relu(self.fc1(x)))


class Net(torch.nn.Module):

    def __init__(self, n=2):
        super().__init__()
        self.emb = torch.nn.Embedding(50, 32)
        self.layer = torch.nn.ModuleList([Block() for _ in range(n)])

    def forward(self, ids):
        x = self.emb(ids)
        for layer in self.layer:
            x = layer(x)
        return x


CFG = {"compression_training": {
    "weight_quantization": {
        "shared_parameters": {"enabled": True, "quantize_weight_in_forward": True, "quantize_groups": 4,
                              "schedule_offset": 1},
        "different_groups": {"wq1": {"params": {"start_bits": 8, "target_bits": 4},
                                     "modules": ["attn.q", "attn.k", "emb"]}}},
    "activation_quantization": {
        "shared_parameters": {"enabled": True, "schedule_offset": 0, "range_calibration": "static"},
        "different_groups": {"aq1": {"params": {"bits": 8}, "modules": ["fc2"]}}},
    "sparse_pruning": {
        "shared_parameters": {"enabled": True, "method": "l1", "schedule_offset": 0},
        "different_groups": {"sp1": {"params": {"dense_ratio": 0.5}, "modules": ["attn.v"]}}},
    "row_pruning": {
        "shared_parameters": {"enabled": True, "method": "topk", "schedule_offset": 0},
        "different_groups": {"rp1": {"params": {"dense_ratio": 0.5}, "modules": ["fc1"],
                                     "related_modules": [["fc2"]]}}},
    "head_pruning": {
        "shared_parameters": {"enabled": True, "method": "topk", "schedule_offset": 0, "num_heads": 4},
        "different_groups": {"hp1": {"params": {"dense_ratio": 0.5}, "modules": ["attn.o"],
                                     "related_modules": [["attn.q", "attn.k", "attn.v"]]}}},
}}


def test_quantizers_and_binarizer():
    torch.manual_seed(0)
    x = torch.randn(8, 64, requires_grad=True)
    for q, bits, levels in ((SymQuantizer, 4, 16), (AsymQuantizer, 4, 17), (TernaryQuantizer, 2, 3),
                            (BinaryQuantizer, 1, 2)):
        y = q.apply(x, bits, None, None, 8)
        assert all(torch.unique(row).numel() <= levels for row in y.detach())
        g, = torch.autograd.grad(y.sum(), x)
        assert torch.equal(g, torch.ones_like(x))  # straight-through
    y8 = SymQuantizer.apply(x, 8, None, None, 8)
    assert (y8 - x).abs().max() <= x.abs().max() / 127
    scores = torch.randn(10, 10, requires_grad=True)
    mask = TopKBinarizer.apply(scores, 0.3, False)
    assert mask.sum() == 30 and scores.detach()[mask.bool()].min() >= scores.detach()[~mask.bool()].max()
    (mask * torch.arange(100.0).view(10, 10)).sum().backward()
    assert torch.equal(scores.grad, torch.arange(100.0).view(10, 10))


def test_init_schedule_train_clean():
    torch.manual_seed(0)
    m = Net()
    dense = Net()
    dense.load_state_dict(m.state_dict())
    init_compression(m, CFG)
    assert isinstance(m.layer[0].fc1, LinearLayer_Compress) and isinstance(m.emb, Embedding_Compress)
    ids = torch.randint(0, 50, (3, 7))
    # nothing enabled yet: identical to the dense model
    torch.testing.assert_close(m(ids), dense(ids))
    sch = compression_scheduler(m, get_compression_config(CFG))
    sch.step(step_zero_check=True)
    assert m.layer[0].fc1.row_pruning_enabled and not m.layer[0].attn.q.weight_quantization_enabled
    sch.step()
    assert m.layer[0].attn.q.weight_quantization_enabled and sch.weight_quantization_enabled
    out = m(ids)
    assert not torch.allclose(out, dense(ids))
    out.sum().backward()
    assert m.layer[0].fc1.row_mask_scores.grad is not None
    assert m.layer[0].attn.o.head_pruning_scores.grad is not None
    assert m.layer[0].attn.q.weight.grad is not None  # STE through the weight quantizer
    m.eval()
    before = m(ids)
    redundancy_clean(m, CFG)
    after = m(ids)
    torch.testing.assert_close(before, after, rtol=1e-5, atol=1e-5)
    b = m.layer[0]
    assert b.fc1.weight.shape == (64, 32) and b.fc2.weight.shape == (32, 64)
    assert b.attn.q.weight.shape == (16, 32) and b.attn.o.weight.shape == (32, 16)
    assert (b.attn.v.weight == 0).float().mean() >= 0.49  # sparse mask baked in (then head rows dropped)
    assert torch.unique(b.attn.q.weight.reshape(4, -1)[0]).numel() <= 16  # 4-bit weights baked in


def test_channel_pruning_conv_bn():
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Conv2d(3, 8, 3, padding=1), torch.nn.BatchNorm2d(8), torch.nn.ReLU(),
                              torch.nn.Conv2d(8, 4, 3, padding=1))
    cfg = {"compression_training": {"channel_pruning": {
        "shared_parameters": {"enabled": True, "method": "l1", "schedule_offset": 0},
        "different_groups": {"cp1": {"params": {"dense_ratio": 0.5}, "modules": ["^0$"],
                                     "related_modules": [["^1$", "^3$"]]}}}}}
    init_compression(net, cfg)
    assert isinstance(net[0], Conv2dLayer_Compress) and isinstance(net[1], BNLayer_Compress)
    compression_scheduler(net, get_compression_config(cfg)).step(step_zero_check=True)
    assert net[0].channel_pruning_enabled
    net.eval()
    x = torch.randn(2, 3, 6, 6)
    before = net(x)
    redundancy_clean(net, cfg)
    assert net[0].weight.shape[0] == 4 and net[1].num_features == 4 and net[3].weight.shape[1] == 4
    torch.testing.assert_close(net(x), before, rtol=1e-5, atol=1e-5)


def test_layer_reduction_student_init():
    torch.manual_seed(0)
    teacher, student = Net(4), Net(2)
    cfg = {"compression_training": {"layer_reduction": {
        "enabled": True, "keep_number_layer": 2, "module_name_prefix": "layer", "teacher_layer": [1, 3],
        "other_module_name": ["emb"]}}}
    init_compression(student, cfg, teacher_model=teacher)
    torch.testing.assert_close(student.layer[0].fc1.weight, teacher.layer[1].fc1.weight)
    torch.testing.assert_close(student.layer[1].attn.o.bias, teacher.layer[3].attn.o.bias)
    torch.testing.assert_close(student.emb.weight, teacher.emb.weight)


def test_compression_config_getters():
    from deepspeed_b200.compression import config as C
    sec = {"weight_quantization": {"shared_parameters": {"enabled": True, "quantize_groups": 4,
                                                          "fp16_mixed_quantize": {"enabled": True}},
                                   "different_groups": {"g1": {"params": {"start_bits": 8, "target_bits": 4},
                                                               "modules": ["attn"]}}},
           "head_pruning": {"shared_parameters": {"enabled": False}},
           "layer_reduction": {"enabled": True, "keep_number_layer": 4}}
    wq = C.get_weight_quantization(sec)
    assert wq["shared_parameters"]["quantize_groups"] == 4 and wq["shared_parameters"]["fp16_mixed_quantize"] is True
    assert wq["different_groups"]["g1"]["params"]["quantization_period"] == 1
    assert wq["different_groups"]["g1"]["modules"] == ["attn"]
    sh = C.get_weight_quantization_shared_parameters(sec["weight_quantization"])
    assert sh["enabled"] is True and sh["quantization_type"] == "symmetric"
    assert list(C.get_weight_quantization_different_groups(sec["weight_quantization"])) == ["g1"]
    assert C.get_sparse_pruning(sec)["shared_parameters"]["enabled"] is False
    assert C.get_row_pruning_shared_parameters({})["method"] == "l1"
    assert C.get_head_pruning_different_groups({}) == {} and C.get_channel_pruning({})["different_groups"] == {}
    assert C.get_activation_quantization_shared_parameters({})["range_calibration"] == "dynamic"
    assert C.get_layer_reduction_enabled(sec) and C.get_layer_reduction_params(sec) == {"keep_number_layer": 4}
    assert C.get_layer_reduction_params({}) is False
    assert C.get_quantize_enabled({"compression_training": sec}) and not C.get_quantize_enabled({})
