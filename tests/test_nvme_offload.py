"""ZeRO-Infinity: optimizer moments swapped to NVMe files through the native aio library must train exactly like the
CPU-offload path; swapper unit tests. Mirrors reference tests/unit/runtime/zero/test_nvme_checkpointing.py and
tests/unit/ops/aio."""
import glob
import os

import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def _train(device, nvme_path, steps=5, save_dir=None, load_dir=None, skip=0):
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = SimpleModel(hidden_dim=32, nlayers=3)
    off = {"device": device, "pin_memory": False}
    if device == "nvme":
        off.update(nvme_path=nvme_path, buffer_count=3)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 4,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-2, "weight_decay": 0.01}},
        "zero_optimization": {"stage": 2, "offload_optimizer": off, "reduce_bucket_size": 2000},
        "aio": {"block_size": 65536, "queue_depth": 4}})
    if load_dir:
        eng.load_checkpoint(load_dir)
    losses = []
    for x, y in random_batches(skip + steps, 4, 32, "cpu", seed=1)[skip:]:
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    if save_dir:
        eng.save_checkpoint(save_dir)
    return losses


def test_nvme_offload_matches_cpu_offload(tmp_path):
    nv = str(tmp_path / "nvme")
    a = run_distributed(_train, 1, args=("cpu", None))[0]
    b = run_distributed(_train, 1, args=("nvme", nv))[0]
    assert a == pytest.approx(b, abs=1e-6), (a, b)
    files = glob.glob(os.path.join(nv, "zero_stage_3", "optimizer", "rank0", "*.tensor.swp"))
    assert len(files) >= 2 and all(os.path.getsize(f) > 0 for f in files)
    assert a[-1] < a[0]


def test_nvme_offload_checkpoint_resume(tmp_path):
    nv, ck = str(tmp_path / "nvme"), str(tmp_path / "ckpt")
    full = run_distributed(_train, 1, args=("nvme", nv, 6))[0]
    run_distributed(_train, 1, args=("nvme", nv + "a", 3, ck))
    # resume in a fresh swap folder: the moments must come from the checkpoint, not from stale swap files
    resumed = run_distributed(_train, 1, args=("nvme", nv + "b", 3, None, ck, 3))[0]
    assert resumed == pytest.approx(full[3:], abs=1e-5), (resumed, full)


def test_pipelined_swapper_roundtrip(tmp_path):
    from deepspeed_b200.runtime.swap_tensor.aio_config import get_aio_config
    from deepspeed_b200.runtime.swap_tensor.pipelined_optimizer_swapper import PipelinedOptimizerSwapper
    sw = PipelinedOptimizerSwapper(None, get_aio_config({"aio": {"block_size": 4096}}), str(tmp_path), 5000, 2,
                                   sets_in_flight=3)
    data = {}
    for k in range(5):
        n = 1000 * (k + 1)
        data[k] = [torch.randn(n), torch.randn(n)]
        sw.register(f"k{k}", n, data[k])
    sw.prefetch("k0")
    for k in range(5):
        bufs = sw.acquire(f"k{k}")
        if k + 1 < 5:
            sw.prefetch(f"k{k + 1}")
        for b, want in zip(bufs, data[k]):
            assert torch.equal(b, want)
            b.add_(1.0)
        sw.release(f"k{k}")
    sw.flush()
    for k in range(5):
        got = sw.read_all(f"k{k}")
        for g, want in zip(got, data[k]):
            assert torch.equal(g, want + 1.0)


def test_param_swapper_and_async_swapper(tmp_path):
    from types import SimpleNamespace
    from deepspeed_b200.ops.aio import aio_handle
    from deepspeed_b200.runtime.swap_tensor.async_swapper import AsyncTensorSwapper, make_staging
    from deepspeed_b200.runtime.swap_tensor.partitioned_param_swapper import (AsyncPartitionedParameterSwapper,
                                                                               PartitionedParamStatus)
    cfg = SimpleNamespace(zero_config=SimpleNamespace(offload_param=SimpleNamespace(
        nvme_path=str(tmp_path), buffer_size=4096, buffer_count=2)), _param_dict={"aio": {"block_size": 1024}})
    sw = AsyncPartitionedParameterSwapper(cfg, torch.bfloat16)
    params = [torch.nn.Parameter(torch.randn(3000).to(torch.bfloat16)) for _ in range(4)]
    keep = [p.detach().clone() for p in params]
    for p in params:
        sw.get_buffer(p, p.numel()).copy_(p.detach())
        sw.swap_out_and_release([p])
    assert sw.available_swap_in_buffers() == 2
    sw.swap_in(params[:2], async_op=True)
    assert sw.status[sw._pid(params[0])] == PartitionedParamStatus.INFLIGHT
    for p, want in zip(params[:2], keep[:2]):
        assert torch.equal(sw.tensor(p), want)
    sw.remove_partition_and_release_buffers(params[:2])
    sw.swap_in(params[2:], async_op=False)
    assert torch.equal(sw.tensor(params[3]), keep[3])
    h = aio_handle(block_size=1024)
    asw = AsyncTensorSwapper(h, 256)
    asw.add_buffers(make_staging(2, 2048, torch.float32))
    ts = [torch.randn(1500) for _ in range(5)]
    paths = [str(tmp_path / f"g{i}.swp") for i in range(5)]
    asw.swap_out_tensors(ts, paths)
    asw.wait()
    for t, path in zip(ts, paths):
        back = torch.empty(1536)
        h.sync_pread(back, path)
        assert torch.equal(back[:1500], t)


def test_nvme_benchmark_engines(tmp_path):
    """The DeepNVMe benchmark engines (reference deepspeed/nvme/*_engine.py, ds_aio_job.py, perf_run_sweep.py)."""
    import sys
    from types import SimpleNamespace
    from deepspeed_b200.nvme.engines import (AIOBasic_Engine, AIOHandle_Engine, Job, SweepConfig, Torch_FastIO_Engine,
                                             TorchIO_Engine, run, run_job)
    args = SimpleNamespace(folder=str(tmp_path), io_size=1 << 20, block_size=1 << 18, queue_depth=4,
                           intra_op_parallelism=2, loops=2, warmup_loops=1, keep_files=True)
    for cls in (AIOBasic_Engine, AIOHandle_Engine, TorchIO_Engine):
        w = run(cls, args, read_op=False)
        r = run(cls, args, read_op=True)
        assert w["op"] == "write" and r["op"] == "read" and w["num_bytes"] == 1 << 20
        assert r["gbps_best"] > 0 and w["loops"] == 2 and r["engine"] == cls.__name__
    w = run(Torch_FastIO_Engine, args, read_op=False)
    assert w["gbps_best"] > 0
    import torch
    f = [p for p in tmp_path.iterdir() if "write" in p.name][0]
    # what the handle engine wrote is what it reads back
    eng = AIOHandle_Engine(args, 7, read_op=False)
    eng.ctxt["buffer"].copy_(torch.arange(1 << 20, dtype=torch.int64).to(torch.uint8))
    eng.write()
    path = eng.ctxt["file"]
    want = eng.ctxt["buffer"].clone()
    eng.fini()
    back = AIOBasic_Engine(SimpleNamespace(**{**vars(args)}), 7, read_op=False)
    back.ctxt["file"] = path
    back.read()
    assert torch.equal(back.ctxt["buffer"], want)
    out = tmp_path / "job.log"
    run_job(Job([sys.executable, "-c", "\"print('hello from job')\""], str(out)))
    assert "hello from job" in out.read_text()
    import pytest
    with pytest.raises(AssertionError):
        run_job(Job([sys.executable, "-c", "\"raise SystemExit(3)\""], str(out)))
    sc = SweepConfig(SimpleNamespace(nvme_dir=[str(tmp_path)], sweep_config={"block_size": ["128K", "1M"],
                                                                             "queue_depth": [4, 8]}, loops=2))
    assert len(sc.combinations()) == 4 and sc.read and sc.write and "--loops 2" in sc.other_options
