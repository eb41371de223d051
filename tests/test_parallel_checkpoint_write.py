"""Data-parallel checkpoint writing (reference runtime/model_checkpointing + io/fast_file_writer.py parallel writers):
stripe writer / reader, writer election, engine save / load at two ranks."""
import os

import pytest
import torch

from common import run_distributed
from deepspeed_b200.io import (Double_IO_Buffer, FastFileWriter, FastFileWriterConfig, MockFileWriter, PyFileWriter,
                               Single_IO_Buffer, StripedReader, find_striped_writers, striped_file_path)
from deepspeed_b200.runtime.model_checkpointing import (CheckpointDataParallel, CheckpointWriterFactory,
                                                        DataParallelWriterFactory, ExpertParallelInfo, MPUInfo,
                                                        UniversalParallelInfo, get_checkpoint_config,
                                                        get_checkpoint_writer_config)


@pytest.mark.parametrize("n_writers,double", [(1, True), (2, True), (3, False)])
def test_striped_write_and_read_roundtrip(tmp_path, n_writers, double):
    torch.manual_seed(0)
    state = {"w": torch.randn(300, 257), "b": torch.arange(1000), "meta": {"step": 7, "name": "x" * 100}}
    path = str(tmp_path / "model_states.pt")
    sizes = []
    for r in range(n_writers):
        fp = striped_file_path(path, r, n_writers) if n_writers > 1 else path
        cfg = FastFileWriterConfig(io_buffer_size=64 * 1024, double_buffer=double, num_parallel_writers=n_writers,
                                   writer_rank=r, stripe_bytes=8192)
        w = FastFileWriter(fp, cfg)
        torch.save(state, w, _use_new_zipfile_serialization=False)
        w.close()
        sizes.append(os.path.getsize(fp))
        assert w.get_stats()["write_bytes"] == w.stream_off and w.get_stats()["close"] == 1
    if n_writers == 1:
        got = torch.load(path, weights_only=False)
    else:
        assert find_striped_writers(path) == n_writers and max(sizes) - min(sizes) <= 8192
        with StripedReader(path, stripe_bytes=8192) as f:
            got = torch.load(f, weights_only=False)
            f.seek(0)
            head = f.read(10)
            f.seek(-5, 2)
            assert len(head) == 10 and len(f.read()) == 5 and f.tell() == sum(sizes)
    assert torch.equal(got["w"], state["w"]) and torch.equal(got["b"], state["b"]) and got["meta"] == state["meta"]
    assert find_striped_writers(str(tmp_path / "other.pt")) is None


def test_io_buffers_fill_and_drain(tmp_path):
    from deepspeed_b200.ops.aio import aio_handle
    h = aio_handle(1 << 16, 4, False, True, 2)
    src = torch.arange(20000, dtype=torch.int32).view(torch.uint8)
    for cls, size in ((Single_IO_Buffer, 8192), (Double_IO_Buffer, 16384)):
        buf = cls(torch.empty(size, dtype=torch.uint8), h)
        path, off, pos = str(tmp_path / cls.__name__), 0, 0
        open(path, "wb").close()
        assert buf.is_empty()
        while pos < src.numel():
            pos += buf.fill(src, pos)
            if buf.is_full():
                n = buf.get_offset()
                assert buf.get_aligned_num_bytes() == n and buf.get_unaligned_num_bytes() == 0
                buf.drain(n, path, off)
                off += n
        n = buf.get_offset()
        buf.drain(n, path, off)
        buf.complete_ongoing_drain()
        assert open(path, "rb").read() == src.numpy().tobytes()
    with pytest.raises(AssertionError):
        Single_IO_Buffer(torch.empty(1000, dtype=torch.uint8), h)


def _info(world, rank, machines, mpu=None, ep=None):
    return UniversalParallelInfo(world, rank, rank % (world // machines), mpu, ep, mpu is None, machines,
                                 rank // (world // machines), machines * 2)


def test_writer_election():
    U = CheckpointDataParallel
    # no unit: only group rank 0 writes model states (every rank still writes its ZeRO shard)
    assert DataParallelWriterFactory(_info(8, 0, 2), None).create_config(0, False).world_size == 1
    assert DataParallelWriterFactory(_info(8, 3, 2), None).create_config(0, False) is None
    assert DataParallelWriterFactory(_info(8, 3, 2), None).create_config(1, False).world_size == 1
    # every replica shares
    c = DataParallelWriterFactory(_info(8, 5, 2), U.REPLICA).create_config(0, False)
    assert (c.world_size, c.rank, c.global_rank, c.pure_dp) == (8, 5, 5, True)
    # one per machine / per socket
    got = [DataParallelWriterFactory(_info(8, r, 2), U.MACHINE).create_config(0, False) for r in range(8)]
    assert [(g.world_size, g.rank) if g else None for g in got] == [(2, 0), None, None, None, (2, 1), None, None, None]
    got = [DataParallelWriterFactory(_info(8, r, 2), U.SOCKET).create_config(0, False) for r in range(8)]
    assert [g.rank if g else None for g in got] == [0, None, 1, None, 2, None, 3, None]
    # ZeRO-3: everything is sharded, every rank writes alone
    assert DataParallelWriterFactory(_info(8, 6, 2), U.MACHINE).create_config(3, False).world_size == 1
    # tp=2 x dp=4 on two machines: the data-parallel peers of lane 1 are 1,3,5,7 -> writers 1 and 5
    mpu = MPUInfo(1, 0, 2, 1, 4, [1, 3, 5, 7], 2)
    c = DataParallelWriterFactory(_info(8, 5, 2, mpu=mpu), U.MACHINE).create_config(0, False)
    assert (c.world_size, c.rank, c.pure_dp) == (2, 1, False)
    assert DataParallelWriterFactory(_info(8, 3, 2, mpu=MPUInfo(1, 0, 2, 1, 4, [1, 3, 5, 7], 1)),
                                     U.MACHINE).create_config(0, False) is None
    # experts: the expert-data-parallel peers decide
    ep = ExpertParallelInfo(4, 2, 2, [2, 6], 1)
    c = DataParallelWriterFactory(_info(8, 6, 2, ep=ep), U.REPLICA).create_config(0, True)
    assert (c.world_size, c.rank) == (2, 1)


def test_checkpoint_config_dicts_and_writer_factory(tmp_path):
    cfg = get_checkpoint_config({"checkpoint": {"tag_validation": "fail", "writer": {"type": "fast",
                                                                                      "data_parallel": "socket"}}})
    assert cfg["tag_validation"] == "FAIL" and cfg["writer"]["type"] == "FAST"
    assert cfg["writer"]["data_parallel"] == "SOCKET" and cfg["writer"]["io_buffer_double"] is True
    assert get_checkpoint_config({})["writer"] is None
    with pytest.raises(AssertionError):
        get_checkpoint_writer_config({"writer": {"type": "carrier-pigeon"}})
    for kind, cls in (("mock", MockFileWriter), ("python", PyFileWriter), ("fast", FastFileWriter)):
        wc = get_checkpoint_writer_config({"writer": {"type": kind, "io_buffer_size": 65536}})
        f = CheckpointWriterFactory(wc, None, None)
        w = f.create_writer(str(tmp_path / f"{kind}.pt"), optimize_dp_state=False)
        assert isinstance(w, cls)
        with pytest.raises(AssertionError):
            f.create_writer(str(tmp_path / "again.pt"), False)
        torch.save({"a": torch.ones(3)}, w, _use_new_zipfile_serialization=False)
        f.release_writer()
        if kind != "mock":
            assert torch.load(str(tmp_path / f"{kind}.pt"), weights_only=False)["a"].sum() == 3


def _engine_shared_write(tmp, zero_stage):
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(64, 256), torch.nn.ReLU(), torch.nn.Linear(256, 64))
    cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "zero_optimization": {"stage": zero_stage},
           "checkpoint": {"writer": {"type": "fast", "data_parallel": "replica", "io_buffer_size": 65536}}}
    engine, *_ = ds.initialize(model=model, config=cfg)
    x = torch.randn(2, 64, device=engine.device)
    for _ in range(2):
        engine.backward(engine(x).square().mean())
        engine.step()
    dp = engine.checkpoint_engine.dp_writer_config
    assert (dp.world_size, dp.rank) == (2, ds.comm.get_rank())
    engine.save_checkpoint(tmp, tag="t")
    files = sorted(os.listdir(os.path.join(tmp, "t")))
    assert "mp_rank_00_model_states.pt-0.2" in files, files
    assert "mp_rank_00_model_states.pt-0.2" in files and "mp_rank_00_model_states.pt-1.2" in files, files
    assert "mp_rank_00_model_states.pt" not in files
    want = {k: v.clone() for k, v in engine.module.state_dict().items()}
    for p in engine.module.parameters():
        p.data.zero_()
    path, _ = engine.load_checkpoint(tmp, tag="t")
    assert path is not None
    for k, v in engine.module.state_dict().items():
        torch.testing.assert_close(v, want[k])
    return files


@pytest.mark.parametrize("zero_stage", [0, 1])
def test_engine_shares_model_state_write_between_replicas(tmp_path, zero_stage):
    out = run_distributed(_engine_shared_write, 2, args=(str(tmp_path), zero_stage))
    if zero_stage == 1:
        assert sum("optim_states" in f for f in out[0]) == 2  # per-rank shards stay whole files


def _engine_decoupled(tmp, fast):
    """Decoupled writer process (plain and through the fast writer with a shared data-parallel write): training goes
    on while the files are written, the checkpoint is committed at a later step and loads back."""
    import deepspeed_b200 as ds
    torch.manual_seed(0)
    model = torch.nn.Sequential(torch.nn.Linear(32, 64), torch.nn.ReLU(), torch.nn.Linear(64, 32))
    writer = {"type": "fast", "decoupled": True, "data_parallel": "replica", "io_buffer_size": 65536} if fast else \
        {"type": "python", "decoupled": True}
    cfg = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
           "zero_optimization": {"stage": 1}, "checkpoint": {"writer": writer}}
    engine, *_ = ds.initialize(model=model, config=cfg)
    assert engine.checkpoint_engine.is_decoupled()
    x = torch.randn(2, 32, device=engine.device)
    engine.backward(engine(x).square().mean())
    engine.step()
    want = {k: v.clone() for k, v in engine.module.state_dict().items()}
    engine.save_checkpoint(tmp, tag="d")
    engine.backward(engine(x).square().mean())
    engine.step()  # the commit is confirmed at a step boundary
    engine.checkpoint_engine.commit_pending(block=True)
    ds.comm.barrier()
    files = sorted(os.listdir(os.path.join(tmp, "d")))
    if fast:
        assert "mp_rank_00_model_states.pt-0.2" in files and "mp_rank_00_model_states.pt-1.2" in files, files
    else:
        assert "mp_rank_00_model_states.pt" in files, files
    assert open(os.path.join(tmp, "latest")).read() == "d"
    engine.load_checkpoint(tmp, tag="d")
    for k, v in engine.module.state_dict().items():
        torch.testing.assert_close(v, want[k])
    engine.destroy()
    return True


@pytest.mark.parametrize("fast", [False, True])
def test_decoupled_writer_process(tmp_path, fast):
    assert all(run_distributed(_engine_decoupled, 2, args=(str(tmp_path), fast)))


def test_checkpoint_size_and_events():
    from deepspeed_b200.runtime.checkpoint_engine import CheckpointSize, DecoupledEvent
    s = CheckpointSize()
    s.set_pre_size(1 << 30)
    s.set_post_size(3 << 30)
    assert s.gb_size() == 2.0 and DecoupledEvent.COMMIT_EVENT.value == 2
