"""``deepspeed_b200.comm`` -- torch.distributed-shaped front-end.

One backend only: ``torch.distributed`` with NCCL on B200 (NVLink 5 / NVSwitch) and gloo for the
CPU plumbing configuration. Every collective is wrapped by :func:`timed_op` so the comms logger can
attribute latency / algbw / busbw per (op, message size).

Reference behaviour studied: ``deepspeed/comm/comm.py:64-760`` (module-level API, ``timed_op``
:106-139, ``init_distributed`` :792-858, ``log_summary`` :439) and ``deepspeed/comm/torch.py:98-466``.
The fused in-kernel collectives live in :mod:`deepspeed_b200.comm.symm`; this module is the baseline
path, the numerical oracle and the fallback for odd shapes/groups.
"""
import datetime
import functools
import os
import time

import torch
import torch.distributed as dist

from ..utils.logging import logger
from .logger import CommsLogger

ReduceOp = dist.ReduceOp
ProcessGroup = dist.ProcessGroup

DEFAULT_TIMEOUT_MIN = int(os.environ.get("DEEPSPEED_TIMEOUT", 30))
default_pg_timeout = datetime.timedelta(minutes=DEFAULT_TIMEOUT_MIN)

comms_logger = CommsLogger()
_initialized_here = False
# debugging switches (reference comm/torch.py:60-90): run with collectives disabled for profiling
_COMM_OFF = {"all": False, "all_reduce": False, "all_gather": False, "reduce_scatter": False, "all_to_all": False}


def set_comm_off(flag=True, ops=("all", )):
    for o in ops:
        _COMM_OFF[o] = flag


def _off(kind):
    return _COMM_OFF["all"] or _COMM_OFF.get(kind, False)


class _NoopWork:
    def wait(self):
        return True

    def is_completed(self):
        return True


# ---------------------------------------------------------------------------------------------
# init / topology
# ---------------------------------------------------------------------------------------------
def is_initialized():
    return dist.is_available() and dist.is_initialized()


def is_available():
    return dist.is_available()


def _device_backend():
    return "nccl" if torch.cuda.is_available() else "gloo"


def mpi_discovery(distributed_port=29500, verbose=True):
    """Populate RANK/WORLD_SIZE/... from an MPI launch (OMPI/PMI env; no mpi4py dependency)."""
    env = os.environ
    for rk, ws, lr in (("OMPI_COMM_WORLD_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_RANK"),
                       ("PMI_RANK", "PMI_SIZE", "MPI_LOCALRANKID"), ("SLURM_PROCID", "SLURM_NTASKS",
                                                                     "SLURM_LOCALID")):
        if rk in env and ws in env:
            env.setdefault("RANK", env[rk])
            env.setdefault("WORLD_SIZE", env[ws])
            env.setdefault("LOCAL_RANK", env.get(lr, "0"))
            env.setdefault("MASTER_ADDR", env.get("MASTER_ADDR", "127.0.0.1"))
            env.setdefault("MASTER_PORT", str(distributed_port))
            if verbose:
                logger.info(f"Discovered launcher env via {rk}: rank={env['RANK']} world={env['WORLD_SIZE']}")
            return True
    return False


def init_distributed(dist_backend=None,
                     auto_mpi_discovery=True,
                     distributed_port=29500,
                     verbose=True,
                     timeout=default_pg_timeout,
                     init_method=None,
                     dist_init_required=None,
                     config=None,
                     rank=-1,
                     world_size=-1):
    """Initialise ``torch.distributed`` (idempotent). Signature mirrors reference ``comm.py:792``."""
    global _initialized_here
    configure(deepspeed_config=config)
    if dist_init_required is False:
        return
    if is_initialized():
        return
    required = ("RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT")
    if init_method is None and rank < 0 and not all(k in os.environ for k in required):
        cloud = False
        if auto_mpi_discovery and "OMPI_COMM_WORLD_RANK" in os.environ:
            if in_aml() and not in_dlts():
                patch_aml_env_for_torch_nccl_backend(verbose=verbose)
                cloud = True
            elif in_aws_sm():
                patch_aws_sm_env_for_torch_nccl_backend(verbose=verbose)
                cloud = True
        if not cloud and not (auto_mpi_discovery and mpi_discovery(distributed_port, verbose)):
            # single-process default: a 1-rank world on localhost
            os.environ.setdefault("RANK", "0")
            os.environ.setdefault("WORLD_SIZE", "1")
            os.environ.setdefault("LOCAL_RANK", "0")
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", str(distributed_port))
    backend = dist_backend or _device_backend()
    kwargs = dict(backend=backend, timeout=timeout)
    if init_method is not None:
        kwargs["init_method"] = init_method
    if rank >= 0:
        kwargs.update(rank=rank, world_size=world_size)
    if backend == "nccl" and torch.cuda.is_available():
        local_rank = int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(local_rank % max(1, torch.cuda.device_count()))
        kwargs["device_id"] = torch.device("cuda", torch.cuda.current_device())
    if verbose and int(os.environ.get("RANK", max(rank, 0))) == 0:
        logger.info(f"Initializing torch.distributed backend={backend}")
    dist.init_process_group(**kwargs)
    _initialized_here = True


def in_aml():
    """Running inside Azure Machine Learning?"""
    return "AZUREML_EXPERIMENT_ID" in os.environ


def in_aws_sm():
    """Running inside AWS SageMaker?"""
    return "SM_TRAINING_ENV" in os.environ


def in_dlts():
    """Running on a DLTS cluster?"""
    return "DLTS_JOB_ID" in os.environ


def patch_aml_env_for_torch_nccl_backend(master_port=6105, verbose=True):
    """Translate the OpenMPI variables Azure ML sets into RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* (reference
    ``comm.py:921``). Multi-node jobs take the master from AZ_BATCH_MASTER_NODE, single-node ones from
    AZ_BATCHAI_MPI_MASTER_NODE."""
    env = os.environ
    env["RANK"] = env["OMPI_COMM_WORLD_RANK"]
    env["WORLD_SIZE"] = env["OMPI_COMM_WORLD_SIZE"]
    single_node = int(env["OMPI_COMM_WORLD_LOCAL_SIZE"]) == int(env["WORLD_SIZE"])
    if not single_node:
        env["MASTER_ADDR"] = env["AZ_BATCH_MASTER_NODE"].split(":")[0]
        env.setdefault("MASTER_PORT", str(master_port))
    else:
        env["MASTER_ADDR"] = env["AZ_BATCHAI_MPI_MASTER_NODE"]
        env["MASTER_PORT"] = str(DEFAULT_AML_MASTER_PORT)
    if verbose:
        logger.info(f"NCCL_SOCKET_IFNAME original value = {env.get('NCCL_SOCKET_IFNAME')}")
    env["NCCL_SOCKET_IFNAME"] = DEFAULT_AML_NCCL_SOCKET_IFNAME
    env["LOCAL_RANK"] = env["OMPI_COMM_WORLD_LOCAL_RANK"]
    if verbose:
        logger.info(f"Discovered AzureML settings of world_rank={env['RANK']}, local_rank={env['LOCAL_RANK']}, "
                    f"world_size={env['WORLD_SIZE']}, master_addr={env['MASTER_ADDR']}, "
                    f"master_port={env['MASTER_PORT']}")


def patch_aws_sm_env_for_torch_nccl_backend(verbose=True):
    """SageMaker launches with OpenMPI: copy its rank variables (MASTER_* are already set there)."""
    env = os.environ
    env["RANK"] = env["OMPI_COMM_WORLD_RANK"]
    env["LOCAL_RANK"] = env["OMPI_COMM_WORLD_LOCAL_RANK"]
    env["WORLD_SIZE"] = env["OMPI_COMM_WORLD_SIZE"]
    if verbose:
        logger.info(f"Discovered AWS SageMaker settings of world_rank={env['RANK']}, local_rank={env['LOCAL_RANK']}, "
                    f"world_size={env['WORLD_SIZE']}, master_addr={env.get('MASTER_ADDR')}, "
                    f"master_port={env.get('MASTER_PORT')}")


DEFAULT_AML_MASTER_PORT = "54965"
DEFAULT_AML_NCCL_SOCKET_IFNAME = "^docker0,lo"


cdb = None  # the backend object (reference ``comm.py`` global of the same name), built on demand


def init_deepspeed_backend(ds_backend, timeout, init_method):
    """The reference can swap in its own (non-torch) backends here; every backend of this framework is
    ``torch.distributed`` plus the in-kernel NVLink collectives of ``comm/symm.py``, so there is nothing to build."""
    logger.debug(f"{ds_backend} backend: torch.distributed process groups are used directly")


def set_backend():
    """Reference ``comm.py:203``: select the active backend object; a single backend exists here."""
    global cdb
    if cdb is None and is_initialized():
        from .backend import TorchBackend
        cdb = TorchBackend(_device_backend())
    return _device_backend()


def destroy_process_group(group=None):
    if is_initialized():
        dist.destroy_process_group(group)


def configure(deepspeed_config=None, enabled=None, prof_all=None, prof_ops=None, verbose=None, debug=None):
    """Configure the comms logger from a ds_config (or explicit overrides)."""
    if deepspeed_config is not None:
        cl = getattr(deepspeed_config, "comms_logger", None) or getattr(deepspeed_config, "comms_config", None)
        if cl is not None:
            comms_logger.configure(cl)
    if enabled is not None:
        comms_logger.enabled = enabled
    if prof_all is not None:
        comms_logger.prof_all = prof_all
    if prof_ops is not None:
        comms_logger.prof_ops = prof_ops
    if verbose is not None:
        comms_logger.verbose = verbose
    if debug is not None:
        comms_logger.debug = debug


def get_rank(group=None):
    return dist.get_rank(group) if is_initialized() else 0


def get_world_size(group=None):
    return dist.get_world_size(group) if is_initialized() else 1


def get_local_rank():
    return int(os.environ.get("LOCAL_RANK", 0))


def get_global_rank(group=None, group_rank=0):
    if group is None or not is_initialized():
        return group_rank
    return dist.get_global_rank(group, group_rank)


def get_world_group():
    return dist.group.WORLD if is_initialized() else None


def get_all_ranks_from_group(group=None):
    if not is_initialized():
        return [0]
    return dist.get_process_group_ranks(group or dist.group.WORLD)


def new_group(ranks=None, **kw):
    return dist.new_group(ranks=ranks, **kw)


def get_backend(group=None):
    return dist.get_backend(group)


def initialize_mesh_device(mesh_shape, mesh_dim_names):
    """Reference ``comm.py:765``: a torch DeviceMesh of the given shape / names (e.g. (dp, sp))."""
    assert is_initialized(), "DeepSpeed backend not set, please initialize it using init_process_group()"
    logger.info(f"Initializing mesh device with shape {mesh_shape} and dim names {mesh_dim_names}")
    return init_device_mesh(mesh_shape, mesh_dim_names)


def enable_symm_mem_for_group(group_name):
    """Make torch symmetric memory usable on a named process group (reference ``comm/torch.py:455``)."""
    assert is_initialized(), "DeepSpeed backend not set, please initialize it using init_process_group()"
    try:
        from torch.distributed._symmetric_memory import enable_symm_mem_for_group as _enable
    except Exception as e:  # noqa: BLE001
        raise RuntimeError(f"Backend {get_backend()} does not support symmetric memory initialization") from e
    _enable(group_name)


def has_all_reduce_coalesced():
    return hasattr(dist, "all_reduce_coalesced")


def has_coalescing_manager():
    return hasattr(dist.distributed_c10d, "_coalescing_manager")


def all_gather_coalesced(output_tensors, input_tensors, group=None, async_op=False):
    """``output_tensors[i]`` (list of world tensors, or one flat tensor) receives ``input_tensors[i]`` of every rank;
    the gathers are issued back to back and, with a coalescing manager, as one group launch."""
    assert is_initialized(), "DeepSpeed backend not set, please initialize it using init_process_group()"
    works = []
    for out, inp in zip(output_tensors, input_tensors):
        if isinstance(out, (list, tuple)):
            works.append(dist.all_gather(list(out), inp, group=group, async_op=True))
        else:
            works.append(dist.all_gather_into_tensor(out, inp, group=group, async_op=True))
    if async_op:
        return works
    for w in works:
        w.wait()
    return None


def has_comm_data():
    """Has the comms logger recorded anything yet?"""
    return comms_logger.has_data()


def get_comm_operation_count():
    return comms_logger.get_total_operations()


def get_logged_comm_ops():
    return comms_logger.get_operation_names()


def init_device_mesh(mesh_shape, mesh_dim_names):
    from torch.distributed.device_mesh import init_device_mesh as _idm
    return _idm("cuda" if torch.cuda.is_available() else "cpu", mesh_shape, mesh_dim_names=mesh_dim_names)


def has_all_gather_into_tensor():
    return hasattr(dist, "all_gather_into_tensor")


def has_reduce_scatter_tensor():
    return hasattr(dist, "reduce_scatter_tensor")


# ---------------------------------------------------------------------------------------------
# timed-op wrapper
# ---------------------------------------------------------------------------------------------
def _msg_size(args, kwargs):
    for a in list(args) + list(kwargs.values()):
        if torch.is_tensor(a):
            return a.numel() * a.element_size()
        if isinstance(a, (list, tuple)) and a and torch.is_tensor(a[0]):
            return sum(t.numel() * t.element_size() for t in a)
    return 0


def timed_op(func):
    """Record latency & size of a collective when the comms logger profiles it."""
    name = func.__name__

    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        prof = kwargs.pop("prof", False)
        log_name = kwargs.pop("log_name", name)
        cl = comms_logger
        if not (cl.enabled and (cl.prof_all or prof or log_name in cl.prof_ops)):
            return func(*args, **kwargs)
        cuda = torch.cuda.is_available()
        if cuda:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
        else:
            t0 = time.perf_counter()
        out = func(*args, **kwargs)
        if hasattr(out, "wait") and kwargs.get("async_op", False):
            out.wait()
        if cuda:
            ev1.record()
            ev1.synchronize()
            ms = ev0.elapsed_time(ev1)
        else:
            ms = (time.perf_counter() - t0) * 1e3
        group = kwargs.get("group", None)
        cl.append(name, log_name, ms, _msg_size(args, kwargs), get_world_size(group))
        return out

    return wrapper


# ---------------------------------------------------------------------------------------------
# collectives
# ---------------------------------------------------------------------------------------------
@timed_op
def all_reduce(tensor, op=ReduceOp.SUM, group=None, async_op=False):
    if not is_initialized() or _off("all_reduce"):
        return _NoopWork() if async_op else None
    return dist.all_reduce(tensor, op=op, group=group, async_op=async_op)


@timed_op
def inference_all_reduce(tensor, op=ReduceOp.SUM, group=None):
    """Reference ``comm/torch.py:186``: CPU tensors of a single-host job go through the shared-memory all-reduce
    (``comm/shm.py``, ``csrc/cpu/shm_comm.cpp``); everything else is a plain all-reduce."""
    if is_initialized() and op == ReduceOp.SUM and group is None and tensor.device.type == "cpu":
        from .shm import shm_all_reduce
        if shm_all_reduce(tensor):
            return None
    if not tensor.is_contiguous():  # the backends reduce a contiguous buffer: go through one and copy back
        buf = tensor.contiguous()
        all_reduce.__wrapped__(buf, op=op, group=group)
        tensor.copy_(buf)
        return None
    return all_reduce.__wrapped__(tensor, op=op, group=group)


@timed_op
def all_reduce_coalesced(tensors, op=ReduceOp.SUM, group=None, async_op=False):
    if not is_initialized() or _off("all_reduce"):
        return _NoopWork() if async_op else None
    works = [dist.all_reduce(t, op=op, group=group, async_op=True) for t in tensors]
    if async_op:
        return works[-1]
    for w in works:
        w.wait()


@timed_op
def reduce(tensor, dst, op=ReduceOp.SUM, group=None, async_op=False):
    if not is_initialized():
        return _NoopWork() if async_op else None
    return dist.reduce(tensor, dst=dst, op=op, group=group, async_op=async_op)


@timed_op
def broadcast(tensor, src, group=None, async_op=False):
    if not is_initialized():
        return _NoopWork() if async_op else None
    return dist.broadcast(tensor, src=src, group=group, async_op=async_op)


def broadcast_object_list(object_list, src=0, group=None, device=None):
    if not is_initialized():
        return
    return dist.broadcast_object_list(object_list, src=src, group=group, device=device)


def all_gather_object(object_list, obj, group=None):
    if not is_initialized():
        object_list[0] = obj
        return
    return dist.all_gather_object(object_list, obj, group=group)


@timed_op
def all_gather(tensor_list, tensor, group=None, async_op=False):
    if not is_initialized() or _off("all_gather"):
        if not is_initialized():
            tensor_list[0].copy_(tensor)
        return _NoopWork() if async_op else None
    return dist.all_gather(tensor_list, tensor, group=group, async_op=async_op)


@timed_op
def all_gather_into_tensor(output_tensor, input_tensor, group=None, async_op=False):
    if not is_initialized() or _off("all_gather"):
        if not is_initialized():
            output_tensor.view(-1)[:input_tensor.numel()].copy_(input_tensor.view(-1))
        return _NoopWork() if async_op else None
    return dist.all_gather_into_tensor(output_tensor, input_tensor, group=group, async_op=async_op)


allgather_fn = all_gather_into_tensor


@timed_op
def reduce_scatter_tensor(output_tensor, input_tensor, op=ReduceOp.SUM, group=None, async_op=False):
    if not is_initialized() or _off("reduce_scatter"):
        if not is_initialized():
            output_tensor.view(-1).copy_(input_tensor.view(-1)[:output_tensor.numel()])
        return _NoopWork() if async_op else None
    if op == ReduceOp.AVG and dist.get_backend(group) == "gloo":
        w = dist.reduce_scatter_tensor(output_tensor, input_tensor, op=ReduceOp.SUM, group=group, async_op=False)
        output_tensor.div_(get_world_size(group))
        return _NoopWork() if async_op else None
    return dist.reduce_scatter_tensor(output_tensor, input_tensor, op=op, group=group, async_op=async_op)


reduce_scatter_fn = reduce_scatter_tensor


@timed_op
def reduce_scatter(output, input_list, op=ReduceOp.SUM, group=None, async_op=False):
    if not is_initialized():
        output.copy_(input_list[0])
        return _NoopWork() if async_op else None
    return dist.reduce_scatter(output, input_list, op=op, group=group, async_op=async_op)


@timed_op
def all_to_all_single(output, input, output_split_sizes=None, input_split_sizes=None, group=None, async_op=False):
    if not is_initialized() or _off("all_to_all"):
        if not is_initialized():
            output.copy_(input)
        return _NoopWork() if async_op else None
    if dist.get_backend(group) == "gloo":
        return _gloo_all_to_all_single(output, input, output_split_sizes, input_split_sizes, group, async_op)
    return dist.all_to_all_single(output,
                                  input,
                                  output_split_sizes=output_split_sizes,
                                  input_split_sizes=input_split_sizes,
                                  group=group,
                                  async_op=async_op)


def _gloo_all_to_all_single(output, input, out_splits, in_splits, group, async_op):
    """gloo has no all_to_all: emulate with all_gather of (padded) inputs. CPU plumbing tests only."""
    ws = get_world_size(group)
    rk = get_rank(group)
    inp = input.contiguous()
    if in_splits is None:
        gathered = [torch.empty_like(inp) for _ in range(ws)]
        dist.all_gather(gathered, inp, group=group)
        per = inp.shape[0] // ws
        chunks = [g[rk * per:(rk + 1) * per] for g in gathered]
    else:
        meta = [None] * ws
        dist.all_gather_object(meta, (list(in_splits), inp.shape[0]), group=group)
        maxrows = max(m[1] for m in meta)
        pad = torch.zeros((maxrows, ) + tuple(inp.shape[1:]), dtype=inp.dtype)
        pad[:inp.shape[0]] = inp
        gathered = [torch.empty_like(pad) for _ in range(ws)]
        dist.all_gather(gathered, pad, group=group)
        chunks = []
        for src in range(ws):
            splits = meta[src][0]
            start = sum(splits[:rk])
            chunks.append(gathered[src][start:start + splits[rk]])
    res = torch.cat(chunks, dim=0) if chunks else output
    output.copy_(res.view_as(output))
    return _NoopWork() if async_op else None


@timed_op
def all_to_all(output_tensor_list, input_tensor_list, group=None, async_op=False):
    if not is_initialized():
        output_tensor_list[0].copy_(input_tensor_list[0])
        return _NoopWork() if async_op else None
    if dist.get_backend(group) == "gloo":
        ws, rk = get_world_size(group), get_rank(group)
        for src in range(ws):
            for dst in range(ws):
                if src == rk and dst == rk:
                    output_tensor_list[rk].copy_(input_tensor_list[rk])
                elif src == rk:
                    dist.send(input_tensor_list[dst].contiguous(), get_global_rank(group, dst), group=group)
                elif dst == rk:
                    buf = torch.empty_like(output_tensor_list[src])
                    dist.recv(buf, get_global_rank(group, src), group=group)
                    output_tensor_list[src].copy_(buf)
        return _NoopWork() if async_op else None
    return dist.all_to_all(output_tensor_list, input_tensor_list, group=group, async_op=async_op)


@timed_op
def send(tensor, dst, group=None, tag=0):
    return dist.send(tensor, dst=dst, group=group, tag=tag)


@timed_op
def recv(tensor, src=None, group=None, tag=0):
    return dist.recv(tensor, src=src, group=group, tag=tag)


@timed_op
def isend(tensor, dst, group=None, tag=0):
    return dist.isend(tensor, dst=dst, group=group, tag=tag)


@timed_op
def irecv(tensor, src=None, group=None, tag=0):
    return dist.irecv(tensor, src=src, group=group, tag=tag)


@timed_op
def gather(tensor, gather_list=None, dst=0, group=None, async_op=False):
    return dist.gather(tensor, gather_list=gather_list, dst=dst, group=group, async_op=async_op)


@timed_op
def scatter(tensor, scatter_list=None, src=0, group=None, async_op=False):
    return dist.scatter(tensor, scatter_list=scatter_list, src=src, group=group, async_op=async_op)


@timed_op
def barrier(group=None, async_op=False, device_ids=None):
    if not is_initialized():
        return _NoopWork() if async_op else None
    return dist.barrier(group=group, async_op=async_op)


@timed_op
def monitored_barrier(group=None, timeout=None, wait_all_ranks=False):
    if not is_initialized():
        return
    if dist.get_backend(group) == "gloo":
        return dist.monitored_barrier(group=group, timeout=timeout, wait_all_ranks=wait_all_ranks)
    return dist.barrier(group=group)


def batch_isend_irecv(p2p_op_list):
    return dist.batch_isend_irecv(p2p_op_list)


P2POp = dist.P2POp


def log_summary(show_straggler=False, return_dict=False):
    """Print (rank 0) the per-op latency/bandwidth table; optionally straggler effect."""
    if is_initialized():
        barrier.__wrapped__()
    res = comms_logger.log_all(print_log=(get_rank() == 0), show_straggler=show_straggler,
                               all_reduce_fn=(all_reduce.__wrapped__ if is_initialized() else None))
    if is_initialized():
        barrier.__wrapped__()
    return res if return_dict else None


def reset_log():
    comms_logger.reset()
