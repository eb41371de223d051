"""``deepspeed.checkpointing``: activation checkpointing with partitioning / CPU offload / RNG tracking.

API and semantics follow reference ``runtime/activation_checkpointing/checkpointing.py`` (``CheckpointFunction``
:567, ``non_reentrant_checkpoint`` :795, ``checkpoint`` :1058, ``CudaRNGStatesTracker`` :134, ``configure`` :1162,
``model_parallel_cuda_manual_seed``): Megatron-compatible ``checkpoint(fn, *args)``; saved activations may be
partitioned across model-parallel ranks (all-gathered in backward), moved to pinned host memory, and RNG state
is restored for the recompute.
"""
import contextlib
import copy

import torch

from .. import comm as dist
from ..utils.logging import logger

_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"
PARTITION_ACTIVATIONS = False
CPU_CHECKPOINT = False
CONTIGUOUS_CHECKPOINTING = False
SYNCHRONIZE = False
PROFILE_TIME = False
num_layers = None
mpu = None
mp_rank = None
mp_size = None
mp_group = None
deepspeed_checkpointing_enabled = False
_configured = False


def _cuda():
    return torch.cuda.is_available()


class CudaRNGStatesTracker:
    """Named CUDA RNG states so dropout inside tensor-parallel regions differs per TP rank but is reproducible."""

    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return copy.copy(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception(f"seed {seed} already exists")
        self.seeds_.add(seed)
        if name in self.states_:
            raise Exception(f"cuda rng state {name} already exists")
        if _cuda():
            orig = torch.cuda.get_rng_state()
            torch.cuda.manual_seed(seed)
            self.states_[name] = torch.cuda.get_rng_state()
            torch.cuda.set_rng_state(orig)
        else:
            orig = torch.get_rng_state()
            torch.manual_seed(seed)
            self.states_[name] = torch.get_rng_state()
            torch.set_rng_state(orig)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception(f"cuda rng state {name} is not added")
        get, set_ = (torch.cuda.get_rng_state, torch.cuda.set_rng_state) if _cuda() else (torch.get_rng_state,
                                                                                           torch.set_rng_state)
        orig = get()
        set_(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = get()
            set_(orig)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed):
    """Default state: same across TP ranks; model-parallel state: seed + 2718 + tp_rank."""
    global mpu
    tp_rank = 0
    if mpu is not None and hasattr(mpu, "get_model_parallel_rank"):
        tp_rank = mpu.get_model_parallel_rank()
    else:
        from ..utils import groups
        try:
            tp_rank = groups._get_model_parallel_rank()
        except Exception:
            tp_rank = 0
    model_parallel_seed = seed + 2718 + tp_rank
    _CUDA_RNG_STATE_TRACKER.reset()
    if _cuda():
        torch.cuda.manual_seed(seed)
    else:
        torch.manual_seed(seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, model_parallel_seed)


model_parallel_reconfigure_tp_seed = model_parallel_cuda_manual_seed


def _mp_info():
    global mp_rank, mp_size, mp_group
    if mp_size is None:
        if mpu is not None:
            mp_rank, mp_size, mp_group = (mpu.get_model_parallel_rank(), mpu.get_model_parallel_world_size(),
                                          mpu.get_model_parallel_group())
        else:
            from ..utils import groups
            try:
                mp_rank, mp_size, mp_group = (groups._get_model_parallel_rank(),
                                              groups._get_model_parallel_world_size(),
                                              groups._get_model_parallel_group())
            except Exception:
                mp_rank, mp_size, mp_group = 0, 1, None
    return mp_rank, mp_size, mp_group


def _partition(t):
    """Keep this MP rank's 1/mp slice of a saved activation."""
    rank, size, _ = _mp_info()
    flat = t.detach().contiguous().view(-1)
    per = (flat.numel() + size - 1) // size
    part = flat.new_zeros(per)
    a, b = rank * per, min((rank + 1) * per, flat.numel())
    if b > a:
        part[:b - a].copy_(flat[a:b])
    return part


def _gather(part, shape, numel):
    rank, size, group = _mp_info()
    out = part.new_empty(size * part.numel())
    dist.all_gather_into_tensor(out, part.contiguous(), group=group)
    return out[:numel].view(shape)


class _Saved:
    __slots__ = ("data", "shape", "numel", "device", "partitioned", "requires_grad", "is_tensor")


def _pack(x):
    s = _Saved()
    s.is_tensor = torch.is_tensor(x)
    if not s.is_tensor:
        s.data = x
        return s
    s.shape, s.numel, s.device, s.requires_grad = x.shape, x.numel(), x.device, x.requires_grad
    s.partitioned = PARTITION_ACTIVATIONS and x.is_floating_point() and _mp_info()[1] > 1 and x.numel() > 0
    data = _partition(x) if s.partitioned else x.detach()
    if CPU_CHECKPOINT and data.is_cuda:
        # asynchronous copy into a pooled pinned buffer on a side stream (runtime/offload_activations.py); a fresh
        # pinned allocation per activation would cost more than the copy
        data = _offload_engine().offload(data)
    s.data = data
    return s


_OFFLOAD_ENGINE = None


def _offload_engine():
    global _OFFLOAD_ENGINE
    if _OFFLOAD_ENGINE is None:
        from .offload_activations import ActivationOffloadEngine
        _OFFLOAD_ENGINE = ActivationOffloadEngine(min_offload_bytes=0, keep_last_count=0, max_in_flight=2)
    return _OFFLOAD_ENGINE


def _unpack(s):
    if not s.is_tensor:
        return s.data
    data = s.data
    if not torch.is_tensor(data):  # ticket of the offload engine
        data = _offload_engine().restore(data)
    elif data.device != s.device:
        data = data.to(s.device, non_blocking=True)
    if s.partitioned:
        data = _gather(data, s.shape, s.numel)
    out = data.detach()
    out.requires_grad_(s.requires_grad)
    return out


class CheckpointFunction(torch.autograd.Function):
    """Reentrant checkpoint: forward under no_grad, recompute + backward in ``backward``."""

    @staticmethod
    def forward(ctx, run_function, all_outputs, *args):
        ctx.run_function = run_function
        ctx.fwd_cpu_rng = torch.get_rng_state()
        ctx.fwd_cuda_rng = torch.cuda.get_rng_state() if _cuda() else None
        ctx.fwd_tracker = get_cuda_rng_tracker().get_states()
        ctx.saved = [_pack(a) for a in args]
        if CPU_CHECKPOINT and _cuda() and SYNCHRONIZE:
            torch.cuda.synchronize()
        with torch.no_grad():
            outputs = run_function(*args)
        if torch.is_tensor(outputs):
            non_tensor = []
            all_outputs += [outputs]
            return outputs
        outs = tuple(outputs)
        all_outputs += list(outs)
        ctx.mark_non_differentiable(*[o for o in outs if torch.is_tensor(o) and not o.is_floating_point()])
        return tuple(o for o in outs if torch.is_tensor(o))

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")
        inputs = [_unpack(s) for s in ctx.saved]
        ctx.saved = None
        bwd_cpu = torch.get_rng_state()
        bwd_cuda = torch.cuda.get_rng_state() if _cuda() else None
        bwd_tracker = get_cuda_rng_tracker().get_states()
        torch.set_rng_state(ctx.fwd_cpu_rng)
        if ctx.fwd_cuda_rng is not None:
            torch.cuda.set_rng_state(ctx.fwd_cuda_rng)
        get_cuda_rng_tracker().set_states(ctx.fwd_tracker)
        with torch.enable_grad():
            outputs = ctx.run_function(*inputs)
        torch.set_rng_state(bwd_cpu)
        if bwd_cuda is not None:
            torch.cuda.set_rng_state(bwd_cuda)
        get_cuda_rng_tracker().set_states(bwd_tracker)
        if torch.is_tensor(outputs):
            outputs = (outputs, )
        outs = [o for o in outputs if torch.is_tensor(o)]
        pairs = [(o, g) for o, g in zip(outs, grads) if o.requires_grad and g is not None]
        if pairs:
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        ret = [None, None]
        for inp in inputs:
            ret.append(inp.grad if torch.is_tensor(inp) else None)
        return tuple(ret)


def non_reentrant_checkpoint(function, *args):
    """Saved-tensor-hook based checkpoint (works with ZeRO-3 re-gather; no reentrant autograd call)."""
    return torch.utils.checkpoint.checkpoint(function, *args, use_reentrant=False)


def checkpoint(function, *args):
    """Checkpoint a model or part of the model (Megatron-compatible signature)."""
    all_outputs = []
    CheckpointFunction.apply(function, all_outputs, *args)
    if len(all_outputs) == 1:
        return all_outputs[0]
    return tuple(all_outputs)


def partition_activations_in_checkpoint(partition_activation):
    global PARTITION_ACTIVATIONS
    PARTITION_ACTIVATIONS = partition_activation


def set_num_layers(nlayers):
    global num_layers
    num_layers = nlayers


def reset():
    """Reset contiguous-buffer bookkeeping between iterations (no persistent buffers in this design)."""
    return


def _configure_using_config_file(config, mpu=None):
    global num_layers, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME
    from ..config import DeepSpeedConfig
    cfg = config if hasattr(config, "activation_checkpointing_config") else DeepSpeedConfig(config, mpu=mpu)
    ac = cfg.activation_checkpointing_config
    PARTITION_ACTIVATIONS = ac.partition_activations
    CONTIGUOUS_CHECKPOINTING = ac.contiguous_memory_optimization
    num_layers = ac.number_checkpoints
    CPU_CHECKPOINT = ac.cpu_checkpointing
    SYNCHRONIZE = ac.synchronize_checkpoint_boundary
    PROFILE_TIME = ac.profile


def configure(mpu_, deepspeed_config=None, partition_activations=None, contiguous_checkpointing=None,
              num_checkpoints=None, checkpoint_in_cpu=None, synchronize=None, profile=None):
    global mpu, num_layers, deepspeed_checkpointing_enabled, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, \
        CPU_CHECKPOINT, SYNCHRONIZE, PROFILE_TIME, _configured, mp_size
    _configured = True
    deepspeed_checkpointing_enabled = True
    mpu = mpu_
    mp_size = None
    if deepspeed_config is not None:
        _configure_using_config_file(deepspeed_config, mpu=mpu)
    if partition_activations is not None:
        PARTITION_ACTIVATIONS = partition_activations
    if contiguous_checkpointing is not None:
        CONTIGUOUS_CHECKPOINTING = contiguous_checkpointing
    if num_checkpoints is not None:
        num_layers = num_checkpoints
    if checkpoint_in_cpu is not None:
        CPU_CHECKPOINT = checkpoint_in_cpu
    if synchronize is not None:
        SYNCHRONIZE = synchronize
    if profile is not None:
        PROFILE_TIME = profile
    if CONTIGUOUS_CHECKPOINTING:
        assert PARTITION_ACTIVATIONS, "Contiguous Checkpointing is only available with partitioned activations. " \
            "Set partitioned activations to true in deepspeed config"
        assert num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"


def is_configured():
    return _configured


# ---- helper names of the reference module (``activation_checkpointing/checkpointing.py:78-470``) -------------------
def detach_variable(inputs, device=None):
    """Detached copies (optionally on ``device``) that keep ``requires_grad``: what backward recomputes from."""
    if not isinstance(inputs, tuple):
        raise RuntimeError("Only tuple of tensors is supported. Got Unsupported input type: ", type(inputs).__name__)
    out = []
    for t in inputs:
        if not torch.is_tensor(t):
            out.append(t)
            continue
        x = (t.to(device=device) if device is not None else t).detach()
        x.requires_grad = t.requires_grad
        out.append(x)
    return tuple(out)


def extract_tensors(all_objects):
    """Split a list / tuple into (tensors, non-tensors, is-tensor flags); order preserved within each part."""
    tensors = [v for v in all_objects if torch.is_tensor(v)]
    others = [v for v in all_objects if not torch.is_tensor(v)]
    flags = [torch.is_tensor(v) for v in all_objects]
    if type(all_objects) is tuple:
        return tuple(tensors), tuple(others), tuple(flags)
    return tensors, others, flags


def merge_tensors(tensor_objects, non_tensor_objects, tensor_flags):
    """Inverse of :func:`extract_tensors`."""
    it_t, it_o = iter(tensor_objects), iter(non_tensor_objects)
    return tuple(next(it_t) if flag else next(it_o) for flag in tensor_flags)


def is_activation_to_checkpoint(item):
    """Floating-point tensors with at least one element per model-parallel rank, unless marked ``no_checkpointing``."""
    size = _mp_info()[1] or 1
    return torch.is_tensor(item) and item.is_floating_point() and item.numel() >= size and \
        not getattr(item, "no_checkpointing", False)


def get_partition_size(item):
    size = _mp_info()[1] or 1
    assert item.numel() % size == 0, "Doesn't handle if partition activation if item is not divisible by mp size"
    return item.numel() // size


def get_partition_start(item):
    rank, size, _ = _mp_info()
    return int(item.numel() / (size or 1) * (rank or 0))


def partition_activations(args, cpu_checkpoint=False, contiguous_checkpoint=False):
    """This model-parallel rank's slice of every activation to checkpoint (other arguments unchanged), optionally on
    the host."""
    out = []
    for item in args:
        if not is_activation_to_checkpoint(item):
            out.append(item)
            continue
        n = get_partition_size(item)
        part = item.detach().contiguous().view(-1).narrow(0, get_partition_start(item), n).clone()
        out.append(part.to("cpu") if cpu_checkpoint else part)
    return out


def get_partitioned_activations_for_backward(args, inputs, contiguous_checkpoint=False):
    """Interleave each partitioned activation with its original size (as the reference's saved-tensor list does), so
    that :func:`gather_partitioned_activations` can rebuild the full tensors."""
    out = []
    for arg, inp in zip(args, inputs):
        if not is_activation_to_checkpoint(arg):
            out.append(arg)
            out.append(None)
            continue
        out.append(inp)
        out.append(torch.tensor(arg.size()))
    return out


def gather_partitioned_activations(tensors, device=None):
    """Inverse of the two functions above: all-gather each (slice, size) pair over the model-parallel group."""
    assert len(tensors) % 2 == 0, f"Expected even count of tensors, instead got {len(tensors)}"
    rank, size, group = _mp_info()
    out = []
    for i in range(len(tensors) // 2):
        item, shape = tensors[2 * i], tensors[2 * i + 1]
        if shape is None or not torch.is_tensor(item):
            out.append(item)
            continue
        shape = [int(s) for s in shape.tolist()]
        if device is not None:
            item = item.to(device)
        if group is None or (size or 1) == 1:
            out.append(item.view(shape))
            continue
        full = item.new_empty(item.numel() * size)
        dist.all_gather_into_tensor(full, item.contiguous(), group=group)
        out.append(full.view(shape))
    return tuple(out)


def get_cpu_activations_for_backward(args, inputs):
    """Host copies of the activations to checkpoint (``cpu_checkpointing`` without partitioning)."""
    return [inp.to("cpu") if is_activation_to_checkpoint(arg) else arg for arg, inp in zip(args, inputs)]


get_offloaded_activations_for_backward = get_cpu_activations_for_backward


def restore_offloaded_activations(tensors, device=None):
    dev = device if device is not None else (torch.device("cuda", torch.cuda.current_device())
                                             if torch.cuda.is_available() else torch.device("cpu"))
    return tuple(t.to(dev) if torch.is_tensor(t) and t.device != dev and t.is_floating_point() else t for t in tensors)
