"""Host-side native ops: CPU Adam/Adagrad/Lion vs torch, rollback, async I/O, fast file writer, ZeRO-Offload."""
import os

import pytest
import torch

from simple_model import SimpleModel, random_batches


def test_cpu_adam_matches_torch_adamw():
    from deepspeed_b200.ops.adam import DeepSpeedCPUAdam
    torch.manual_seed(0)
    p = torch.randn(100003)
    ref = p.clone().requires_grad_(True)
    opt = DeepSpeedCPUAdam([p], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    ropt = torch.optim.AdamW([ref], lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    lp = torch.zeros_like(p, dtype=torch.bfloat16)
    for _ in range(4):
        g = torch.randn_like(p)
        opt.step_flat({p: g.to(torch.bfloat16)}, {p: lp}, grad_scale=1.0)
        ref.grad = g.to(torch.bfloat16).float()
        ropt.step()
    torch.testing.assert_close(p, ref.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(lp, p.to(torch.bfloat16), rtol=0, atol=0)


def test_cpu_adam_rollback_restores_state():
    from deepspeed_b200.ops.adam import DeepSpeedCPUAdam
    torch.manual_seed(1)
    p = torch.randn(5000)
    opt = DeepSpeedCPUAdam([p], lr=1e-2, weight_decay=0.01)
    g1, g2 = torch.randn_like(p), torch.randn_like(p)
    opt.step_flat({p: g1})
    snap = (p.clone(), opt.state[p]["exp_avg"].clone(), opt.state[p]["exp_avg_sq"].clone())
    opt.step_flat({p: g2})
    opt.rollback({p: g2})
    torch.testing.assert_close(p, snap[0], rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(opt.state[p]["exp_avg"], snap[1], rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(opt.state[p]["exp_avg_sq"], snap[2], rtol=1e-3, atol=1e-6)


def test_cpu_adagrad_and_lion():
    from deepspeed_b200.ops.adagrad import DeepSpeedCPUAdagrad
    from deepspeed_b200.ops.lion import DeepSpeedCPULion
    from deepspeed_b200.ops.optimizers import FusedAdagrad, FusedLion
    torch.manual_seed(2)
    for native, pyref, kw in ((DeepSpeedCPUAdagrad, FusedAdagrad, dict(lr=1e-2, weight_decay=0.01)),
                              (DeepSpeedCPULion, FusedLion, dict(lr=1e-3, betas=(0.9, 0.99), weight_decay=0.1))):
        a, b = torch.randn(4097), None
        b = a.clone()
        oa, ob = native([a], **kw), pyref([b], **kw)
        for _ in range(3):
            g = torch.randn_like(a)
            oa.step_flat({a: g})
            ob.step_flat({b: g})
        torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)


def test_aio_roundtrip_and_fast_writer(tmp_path):
    from deepspeed_b200.io import FastFileWriter, FastFileWriterConfig
    from deepspeed_b200.ops.aio import aio_handle
    h = aio_handle(block_size=1 << 16, queue_depth=4, intra_op_parallelism=3)
    src = torch.randn(1 << 18)
    path = str(tmp_path / "blob.bin")
    h.async_pwrite(src, path)
    h.wait()
    dst = torch.zeros_like(src)
    h.sync_pread(dst, path)
    assert torch.equal(src, dst)
    # partial read at an offset
    part = torch.zeros(1000)
    h.sync_pread(part, path, file_offset=4 * 123)
    assert torch.equal(part, src[123:1123])
    # torch.save through the double-buffered writer
    p2 = str(tmp_path / "ck.pt")
    w = FastFileWriter(p2, FastFileWriterConfig(io_buffer_size=1 << 16, num_threads=2))
    obj = {"a": torch.arange(100000), "b": [1, 2, 3]}
    torch.save(obj, w, _use_new_zipfile_serialization=False)
    w.close()
    back = torch.load(p2, weights_only=False)
    assert torch.equal(back["a"], obj["a"]) and back["b"] == [1, 2, 3]


@pytest.mark.parametrize("stage", [2, 3])
def test_zero_offload_matches_torch(stage):
    """optimizer state + step on host (DeepSpeedCPUAdam), grads D2H / params H2D (here: all CPU, 1 rank)."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    m = SimpleModel(16, 3)
    ref = SimpleModel(16, 3)
    ref.load_state_dict(m.state_dict())
    eng, _, _, _ = ds.initialize(model=m, config={
        "train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.01}},
        "zero_optimization": {"stage": stage, "offload_optimizer": {"device": "cpu"}, "unit_max_numel": 300,
                              "stage3_param_persistence_threshold": 20}})
    from deepspeed_b200.ops.adam import DeepSpeedCPUAdam
    assert isinstance(eng.basic_optimizer, DeepSpeedCPUAdam)
    ropt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.01)
    for x, y in random_batches(4, 4, 16, "cpu"):
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        ropt.zero_grad()
        ref(x, y).backward()
        ropt.step()
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param
    for (n, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(safe_get_full_fp32_param(p), q.detach(), rtol=1e-4, atol=1e-5)
