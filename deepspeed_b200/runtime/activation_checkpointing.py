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
        host = torch.empty(data.shape, dtype=data.dtype, device="cpu", pin_memory=True)
        host.copy_(data, non_blocking=True)
        data = host
    s.data = data
    return s


def _unpack(s):
    if not s.is_tensor:
        return s.data
    data = s.data
    if data.device != s.device:
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
