

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
