"""Backend objects of the communication layer (reference ``comm/backend.py``, ``comm/torch.py``).

The module-level functions of :mod:`deepspeed_b200.comm.comm` are the implementation; :class:`TorchBackend` is the
object face of the same operations for code written against the reference's ``cdb`` ("communication data backend")
handle: ``backend.all_reduce(t)``, ``backend.get_world_size()``... One backend exists - ``torch.distributed`` process
groups (NCCL on GPUs, gloo on CPU) with the NVLink kernels of ``comm/symm.py`` underneath the fused paths."""
import torch


class Backend:

    def __init__(self, name="backend", rank=0, size=1):
        self.name = name
        self.world_group = None
        self.world_size = size
        self.world_rank = rank
        self.process_groups = []
        self.initialized = False

    def is_initialized(self):
        return self.initialized

    def new_group(self):
        pass

    def init_process_group(self):
        self.initialized = True


class Noop:
    """Finished work handle returned by operations that had nothing to wait for."""

    def wait(self):
        return None

    def is_completed(self):
        return True


_OPS = ("all_reduce", "inference_all_reduce", "all_reduce_coalesced", "reduce", "reduce_scatter", "broadcast",
        "broadcast_object_list", "all_gather", "all_gather_into_tensor", "all_gather_coalesced", "all_gather_object",
        "reduce_scatter_tensor", "all_to_all_single", "all_to_all", "send", "recv", "isend", "irecv", "gather",
        "scatter", "barrier", "monitored_barrier", "get_rank", "get_world_size", "get_backend", "new_group",
        "get_global_rank", "get_world_group", "destroy_process_group", "init_device_mesh",
        "enable_symm_mem_for_group", "has_all_gather_into_tensor", "has_reduce_scatter_tensor")


class TorchBackend(Backend):
    """``torch.distributed`` behind the reference's backend interface. Constructing it initialises the default
    process group when that has not happened yet."""

    def __init__(self, backend=None, timeout=None, init_method=None, rank=-1, world_size=-1, name="torch"):
        super().__init__(name=name)
        from . import comm as C
        self._C = C
        self.using_mpi = False
        if not torch.distributed.is_initialized():
            C.init_distributed(dist_backend=backend, timeout=timeout, init_method=init_method, rank=rank,
                               world_size=world_size, verbose=False)
        self.init_process_group(backend, timeout, init_method, rank, world_size)

    def init_process_group(self, backend=None, timeout=None, init_method=None, rank=-1, world_size=-1):
        self.initialized = torch.distributed.is_initialized()
        if self.initialized:
            self.world_group = torch.distributed.group.WORLD
            self.world_rank = torch.distributed.get_rank()
            self.world_size = torch.distributed.get_world_size()

    def is_initialized(self):
        return torch.distributed.is_initialized()

    def all_gather_base(self, output_tensor, input_tensor, group=None, async_op=False):
        return self._C.all_gather_into_tensor(output_tensor, input_tensor, group=group, async_op=async_op)

    def get_all_gather_function(self):
        return torch.distributed.all_gather_into_tensor

    def get_reduce_scatter_function(self):
        return torch.distributed.reduce_scatter_tensor

    def _reduce_op(self, op):
        return op


def _delegate(name):

    def method(self, *args, **kwargs):
        return getattr(self._C, name)(*args, **kwargs)

    method.__name__ = name
    return method


for _n in _OPS:
    setattr(TorchBackend, _n, _delegate(_n))


# ---- module-level switches and helpers of the reference's ``comm/torch.py`` --------------------------------
def _switch(op):

    def setter(flag=False):
        from . import comm as C
        C.set_comm_off(bool(flag), ops=(op, ))

    setter.__name__ = f"{op}_comm_off"
    setter.__doc__ = f"Turn the ``{op}`` collectives into no-ops (profiling compute without communication)."
    return setter


def all_gather_comm_off(flag=False):
    _switch("all_gather")(flag)


def reduce_scatter_comm_off(flag=False):
    _switch("reduce_scatter")(flag)


def all_reduce_comm_off(flag=False):
    _switch("all_reduce")(flag)


def reduce_comm_off(flag=False):
    _switch("all_reduce")(flag)
    _switch("reduce_scatter")(flag)


def broadcast_comm_off(flag=False):
    _switch("broadcast")(flag)


def backward_comm_off(flag=False):
    """Everything a backward pass communicates: gradient reductions."""
    all_reduce_comm_off(flag)
    reduce_scatter_comm_off(flag)


def disable_compiler_collective(func):
    """The reference keeps its collectives out of ``torch.compile`` graphs with this decorator; nothing here is traced,
    so the function is returned as it is."""
    return func


def build_shm_op():
    """The shared-memory all-reduce for CPU tensors of a single-host job (``comm/shm.py`` over
    ``csrc/cpu/shm_comm.cpp``), or ``None`` when the native library is not built."""
    try:
        from . import shm
        return shm
    except Exception:  # noqa: BLE001
        return None


def get_coalescing_manager(group, device, reqs, async_op=False):
    """``torch.distributed``'s coalescing context (several collectives issued as one)."""
    import torch.distributed.distributed_c10d as c10d
    return c10d._coalescing_manager(group, device=device, async_ops=async_op)
