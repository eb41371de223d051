

def _backend_object():
    import torch
    import deepspeed_b200.comm as dist
    from deepspeed_b200.comm import comm as C
    from deepspeed_b200.comm.backend import Backend, Noop, TorchBackend
    b = TorchBackend("gloo")
    assert isinstance(b, Backend) and b.is_initialized() and b.get_world_size() == 2 == b.world_size
    t = torch.ones(4) * (b.get_rank() + 1)
    b.all_reduce(t)
    assert t.tolist() == [3.0] * 4
    out = torch.empty(8)
    b.all_gather_base(out, torch.full((4, ), float(b.get_rank())))
    assert out.tolist() == [0.0] * 4 + [1.0] * 4
    assert b.has_all_gather_into_tensor() and Noop().wait() is None and Noop().is_completed()
    C.set_backend()
    assert isinstance(C.cdb, TorchBackend)
    return True


def test_backend_object_delegates_to_the_module_functions():
    from common import run_distributed
    assert all(run_distributed(_backend_object, 2))


def test_comms_config_holder():
    from deepspeed_b200.comm.config import DeepSpeedCommsConfig
    c = DeepSpeedCommsConfig({"comms_logger": {"enabled": True, "prof_ops": ["all_reduce"]}})
    assert c.comms_logger_enabled and c.comms_logger.prof_ops == ["all_reduce"]
    assert not DeepSpeedCommsConfig({}).comms_logger_enabled


def test_comm_utils_introspection(monkeypatch):
    import torch
    from deepspeed_b200.comm import comm as C
    from deepspeed_b200.comm import utils as U
    from deepspeed_b200.comm.backend import (all_gather_comm_off, backward_comm_off, build_shm_op,
                                             disable_compiler_collective)
    t = torch.zeros(10, dtype=torch.float16)
    assert U.get_tensor_position(C.all_reduce) == 0 and U.get_tensor_position(C.barrier) == -1
    assert U.get_msg_size_from_args(C.all_reduce, t) == 20
    assert U.get_msg_size_from_args(C.all_reduce, tensor=t) == 20 and U.get_msg_size_from_args(C.barrier) == 0
    assert U.get_msg_size_from_args(C.all_reduce_coalesced, [t, t]) == 40
    assert U.get_default_args(C.all_reduce)["async_op"] is False
    monkeypatch.setenv("OMPI_COMM_WORLD_RANK", "3")
    monkeypatch.delenv("RANK", raising=False)
    assert U.get_world_rank_from_launcher() == 3
    assert U.get_debug_log_name({"log_name": "all_reduce"}, False) == "all_reduce"
    assert U.get_debug_log_name({"log_name": "all_reduce"}, True).startswith("all_reduce | [Caller Func: ")
    all_gather_comm_off(True)
    assert C._off("all_gather")
    all_gather_comm_off(False)
    backward_comm_off(True)
    assert C._off("all_reduce") and C._off("reduce_scatter")
    backward_comm_off(False)
    assert not C._off("all_reduce") and disable_compiler_collective(len) is len and build_shm_op() is not None


def test_io_utils_roundtrip():
    import torch
    from deepspeed_b200.io import bytes_to_tensor, obj_serialization_details, tensor_to_bytes
    t = torch.arange(6, dtype=torch.float32)
    b = tensor_to_bytes(t)
    assert b.dtype == torch.uint8 and b.numel() == 24
    assert torch.equal(bytes_to_tensor(bytes(b.numpy().tobytes()), torch.float32), t)
    d = obj_serialization_details()
    assert (d(t).size, d(t).nbytes, d(t).dtype) == (6, 24, torch.float32)
    st = d(t.untyped_storage())
    assert st.nbytes == 24
