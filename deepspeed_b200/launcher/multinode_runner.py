"""Multi-node command builders (reference ``launcher/multinode_runner.py``: PDSH, OpenMPI, MPICH, IMPI, MVAPICH,
Slurm). Only the command line + exported environment are produced; execution is ``subprocess.Popen`` in the runner."""
import os
import shutil
import sys
from abc import ABC, abstractmethod
from shlex import split

PDSH_MAX_FAN_OUT = 1024


class MultiNodeRunner(ABC):

    def __init__(self, args, world_info_base64):
        self.args = args
        self.validate_args()
        self.user_arguments = list(args.user_args)
        self.user_script = args.user_script
        self.world_info_base64 = world_info_base64
        self.exports = {}

    @abstractmethod
    def backend_exists(self):
        ...

    @abstractmethod
    def get_cmd(self, environment, active_resources):
        ...

    def add_export(self, key, var):
        self.exports[key.strip()] = f"\"{var.strip()}\"" if " " in var.strip() else var.strip()

    def parse_user_args(self):
        return self.args.user_args

    @property
    def name(self):
        return self.__class__.__name__

    def validate_args(self):
        pass

    def _python(self):
        exec_ = []
        if not self.args.no_python:
            exec_ = [sys.executable, "-u"]
            if self.args.module:
                exec_.append("-m")
        return exec_


class PDSHRunner(MultiNodeRunner):

    def backend_exists(self):
        return shutil.which("pdsh") is not None

    @property
    def name(self):
        return "pdsh"

    def get_cmd(self, environment, active_resources):
        environment["PDSH_RCMD_TYPE"] = "ssh"
        if self.args.ssh_port is not None:
            environment["PDSH_SSH_ARGS_APPEND"] = f"{environment.get('PDSH_SSH_ARGS_APPEND', '')} -p {self.args.ssh_port}"
        active_workers = ",".join(active_resources.keys())
        pdsh_cmd = ["pdsh", "-S", "-f", str(PDSH_MAX_FAN_OUT), "-w", active_workers] + split(self.args.launcher_args)
        exports = "".join(f"export {k}={v}; " for k, v in self.exports.items())
        launch = [exports, f"cd {os.path.abspath('.')};", sys.executable, "-u", "-m", "deepspeed_b200.launcher.launch",
                  f"--world_info={self.world_info_base64}", "--node_rank=%n", f"--master_addr={self.args.master_addr}",
                  f"--master_port={self.args.master_port}"]
        for flag in ("no_python", "module", "no_local_rank", "save_pid", "elastic_training"):
            if getattr(self.args, flag):
                launch.append(f"--{flag}")
        kill = pdsh_cmd + ["pkill -f deepspeed_b200.launcher.launch"]
        return pdsh_cmd + launch + [self.user_script] + self.user_arguments, kill, environment


class OpenMPIRunner(MultiNodeRunner):

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool
        self.add_export("UCX_TLS", "tcp")

    def backend_exists(self):
        return shutil.which("ompi_info") is not None

    @property
    def name(self):
        return "openmpi"

    def validate_args(self):
        if self.args.include != "" or self.args.exclude != "":
            raise ValueError(f"{self.name} backend does not support worker include/exclusion")
        if self.args.num_nodes != -1 or self.args.num_gpus != -1:
            raise ValueError(f"{self.name} backend does not support limiting num nodes/gpus")

    def get_cmd(self, environment, active_resources):
        total = sum(len(v) if isinstance(v, (list, tuple)) else v for v in self.resource_pool.values())
        cmd = ["mpirun", "-n", f"{total}", "-hostfile", f"{self.args.hostfile}", "--mca", "btl", "^openib", "--mca",
               "btl_tcp_if_include", "eth0"] + split(self.args.launcher_args)
        export_cmd = []
        for k, v in self.exports.items():
            export_cmd += ["-x", f"{k}={v}"]
        return cmd + export_cmd + self._python() + [self.user_script] + self.user_arguments, None, environment


class MPICHRunner(OpenMPIRunner):

    def backend_exists(self):
        return shutil.which("mpirun") is not None

    @property
    def name(self):
        return "mpich"

    def get_cmd(self, environment, active_resources):
        devices_per_node = list(self.resource_pool.values())
        total = sum(len(v) if isinstance(v, (list, tuple)) else v for v in devices_per_node)
        per = devices_per_node[0] if not isinstance(devices_per_node[0], (list, tuple)) else len(devices_per_node[0])
        cmd = ["mpirun", "-n", f"{total}", "-ppn", f"{per}"] + split(self.args.launcher_args)
        export_cmd = []
        for k, v in self.exports.items():
            export_cmd += ["-genv", f"{k}", f"{v}"]
        export_cmd += ["-genv", "MASTER_ADDR", str(self.args.master_addr), "-genv", "MASTER_PORT",
                       str(self.args.master_port), "-genv", "WORLD_SIZE", str(total), "-genv", "LOCAL_SIZE", str(per)]
        return cmd + export_cmd + self._python() + [self.user_script] + self.user_arguments, None, environment


class IMPIRunner(MPICHRunner):

    @property
    def name(self):
        return "impi"


class MVAPICHRunner(OpenMPIRunner):

    def backend_exists(self):
        return shutil.which("mpiname") is not None

    @property
    def name(self):
        return "mvapich"

    def get_cmd(self, environment, active_resources):
        total = sum(len(v) if isinstance(v, (list, tuple)) else v for v in self.resource_pool.values())
        per = list(self.resource_pool.values())[0]
        per = len(per) if isinstance(per, (list, tuple)) else per
        cmd = ["mpirun", "-np", f"{total}", "-ppn", f"{per}", "--hostfile", f"{self.args.hostfile}"] + \
            split(self.args.launcher_args)
        export_cmd = []
        for k, v in self.exports.items():
            export_cmd += ["-env", f"{k}={v}"]
        return cmd + export_cmd + self._python() + [self.user_script] + self.user_arguments, None, environment


class SlurmRunner(MultiNodeRunner):

    def __init__(self, args, world_info_base64, resource_pool):
        super().__init__(args, world_info_base64)
        self.resource_pool = resource_pool

    def backend_exists(self):
        return shutil.which("sinfo") is not None

    @property
    def name(self):
        return "slurm"

    def get_cmd(self, environment, active_resources):
        total = sum(len(v) if isinstance(v, (list, tuple)) else v for v in self.resource_pool.values())
        cmd = ["srun", "-n", f"{total}"] + split(self.args.launcher_args)
        if getattr(self.args, "comment", ""):
            cmd += ["--comment", self.args.comment]
        if self.args.include != "":
            cmd += ["--include", f"{self.args.include}"]
        if self.args.exclude != "":
            cmd += ["--exclude", f"{self.args.exclude}"]
        if self.args.num_nodes > 0:
            cmd += ["--nodes", f"{self.args.num_nodes}"]
        if self.args.num_gpus > 0:
            cmd += ["--gpus", f"{self.args.num_gpus}"]
        exports = "--export=ALL"
        for k, v in self.exports.items():
            exports += f",{k}={v}"
        return cmd + [exports] + self._python() + [self.user_script] + self.user_arguments, None, environment


def get_runner(name, args, world_info_base64, resource_pool):
    name = name.lower()
    if name == "pdsh":
        return PDSHRunner(args, world_info_base64)
    table = {"openmpi": OpenMPIRunner, "mpich": MPICHRunner, "impi": IMPIRunner, "mvapich": MVAPICHRunner,
             "slurm": SlurmRunner}
    if name not in table:
        raise NotImplementedError(f"Unknown launcher {name}")
    return table[name](args, world_info_base64, resource_pool)
