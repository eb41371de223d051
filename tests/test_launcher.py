"""Launcher: hostfile parsing, include/exclude filters, world-info round trip, generated command lines and the
per-node env (reference tests/unit/launcher/test_run.py, test_multinode_runner.py)."""
import os
import subprocess
import sys

import pytest

from deepspeed_b200.launcher import launch, runner
from deepspeed_b200.launcher.multinode_runner import get_runner

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_hostfile_and_filters():
    pool = runner._parse_hostfile(["worker-0 slots=4\n", "# c\n", "worker-1 slots=4\n"])
    assert pool == {"worker-0": 4, "worker-1": 4}
    act = runner.parse_inclusion_exclusion(pool, "worker-0@worker-1:0,2", "")
    assert act == {"worker-0": [0, 1, 2, 3], "worker-1": [0, 2]}
    act = runner.parse_inclusion_exclusion(pool, "", "worker-1:0")
    assert act["worker-1"] == [1, 2, 3]
    with pytest.raises(ValueError):
        runner.parse_inclusion_exclusion(pool, "worker-9", "")
    with pytest.raises(ValueError):
        runner._parse_hostfile(["worker-0 slots=4\n", "worker-0 slots=2\n"])


def test_world_info_and_rank_env():
    info = {"a": [0, 1], "b": [0, 1, 2]}
    assert runner.decode_world_info(runner.encode_world_info(info)) == info
    envs, world = launch.build_rank_env(info, 1, "a", 29500)
    assert world == 5 and [e["RANK"] for e in envs] == ["2", "3", "4"] and envs[0]["LOCAL_SIZE"] == "3"
    assert envs[0]["CROSS_RANK"] == "1" and envs[0]["MASTER_ADDR"] == "a"


def test_multinode_command_lines():
    args = runner.parse_args(["--launcher", "pdsh", "--master_addr", "w0", "train.py", "--x", "1"])
    r = get_runner("pdsh", args, runner.encode_world_info({"w0": [0], "w1": [0]}), {"w0": 1, "w1": 1})
    cmd, kill, env = r.get_cmd({}, {"w0": [0], "w1": [0]})
    assert cmd[0] == "pdsh" and "w0,w1" in cmd and "train.py" in cmd and env["PDSH_RCMD_TYPE"] == "ssh"
    args = runner.parse_args(["--launcher", "slurm", "--num_nodes", "2", "train.py"])
    r = get_runner("slurm", args, "x", {"w0": 8, "w1": 8})
    cmd, _, _ = r.get_cmd({}, {})
    assert cmd[:3] == ["srun", "-n", "16"] and "--nodes" in cmd
    args = runner.parse_args(["--launcher", "openmpi", "train.py"])
    r = get_runner("openmpi", args, "x", {"w0": 8})
    cmd, _, _ = r.get_cmd({}, {})
    assert cmd[0] == "mpirun" and "-x" in cmd


def test_single_node_launch_runs_and_propagates_failure(tmp_path):
    ok = tmp_path / "ok.py"
    ok.write_text("import os\nopen(os.environ['OUT'] + os.environ['RANK'], 'w').write(os.environ['WORLD_SIZE'])\n")
    env = dict(os.environ, OUT=str(tmp_path / "r"), PYTHONPATH=REPO)
    wi = runner.encode_world_info({"localhost": [0, 1]})
    subprocess.check_call([sys.executable, "-m", "deepspeed_b200.launcher.launch", f"--world_info={wi}", str(ok)], env=env)
    assert (tmp_path / "r0").read_text() == "2" and (tmp_path / "r1").read_text() == "2"
    bad = tmp_path / "bad.py"
    bad.write_text("import os, sys, time\nif os.environ['RANK']=='1': sys.exit(3)\ntime.sleep(30)\n")
    rc = subprocess.call([sys.executable, "-m", "deepspeed_b200.launcher.launch", f"--world_info={wi}", str(bad)], env=env)
    assert rc == 3


def test_elasticity_config_object_and_logger_factory():
    import logging
    import pytest
    from deepspeed_b200.elasticity.config import ElasticityConfig, ElasticityConfigError
    from deepspeed_b200.utils.logging import LoggerFactory, get_log_level_from_string, should_log_le
    c = ElasticityConfig({"enabled": True, "max_train_batch_size": 2000, "micro_batch_sizes": [2, 4, 6], "min_gpus": 2})
    assert (c.max_acceptable_batch_size, c.micro_batches, c.min_gpus, c.max_gpus, c.version) == (2000, [2, 4, 6], 2,
                                                                                                   10000, 0.2)
    assert '"min_gpus": 2' in repr(c) and c.repr()["enabled"] is True
    assert ElasticityConfig({}).micro_batches == [2, 4, 6]
    for bad in ({"enabled": True}, {"enabled": True, "max_train_batch_size": 8},
                {"micro_batch_sizes": 4}, {"micro_batch_sizes": [2, -1]}, {"micro_batch_sizes": [2.5]},
                {"min_gpus": 0}, {"min_gpus": 8, "max_gpus": 4}, {"model_parallel_size": 0}, {"min_time": -1},
                {"num_gpus_per_node": 0}):
        with pytest.raises(ElasticityConfigError):
            ElasticityConfig(bad)
    lg = LoggerFactory.create_logger("dsb200-test", logging.ERROR)
    assert lg.level == logging.ERROR and lg.handlers and not lg.propagate
    with pytest.raises(ValueError):
        LoggerFactory.create_logger()
    assert get_log_level_from_string("Warning") == logging.WARNING and isinstance(should_log_le("debug"), bool)
    with pytest.raises(ValueError):
        get_log_level_from_string("loud")


def test_comet_events_are_rate_limited():
    from deepspeed_b200.monitor.monitor import CometMonitor, EventsLogScheduler
    s = EventsLogScheduler(100)
    assert [s.needs_logging("loss", x) for x in (0, 50, 100, 150, 220)] == [True, False, True, False, True]
    assert s.needs_logging("lr", 10)

    class Exp:
        logged = []

        def log_metric(self, name, value, step=None):
            Exp.logged.append((name, step))

    class Cfg:
        enabled, samples_log_interval = False, 10

    m = CometMonitor(Cfg())
    m.enabled, m._exp = True, Exp()
    m.write_events([("loss", 1.0, 0), ("loss", 1.0, 5), ("loss", 1.0, 12)])
    assert Exp.logged == [("loss", 0), ("loss", 12)]
