"""Token drop / gather across the tensor-parallel ranks around the expert all-to-all.

Reference: ``moe/mappings.py`` (``drop_tokens`` / ``gather_tokens`` and their autograd functions, after Megatron-LM's
``mappings.py``). With tensor parallelism the TP ranks of a replica hold identical activations; when the experts
themselves are not tensor-parallel, sending all of them through the expert all-to-all would do the same expert work
``tp`` times. ``drop_tokens`` keeps this rank's 1/tp slice along ``dim`` before the dispatch, ``gather_tokens``
all-gathers the slices afterwards; each is the other's backward."""
import torch

from .. import comm as dist
from ..utils import groups


def _tp():
    try:
        world = groups._get_model_parallel_world_size()
    except Exception:  # noqa: BLE001  (groups not initialised: no tensor parallelism)
        return 1, 0, None
    if world == 1:
        return 1, 0, None
    return world, groups._get_model_parallel_rank(), groups._get_model_parallel_group()


def _gather_tokens(input_, dim=0):
    world, _, group = _tp()
    if world == 1:
        return input_
    x = input_.contiguous()
    if dim != 0:
        x = x.movedim(dim, 0).contiguous()
    out = x.new_empty((world * x.shape[0], ) + tuple(x.shape[1:]))
    dist.all_gather_into_tensor(out, x, group=group)
    return out if dim == 0 else out.movedim(0, dim).contiguous()


def _drop_tokens(input_, dim=0):
    world, rank, _ = _tp()
    if world == 1:
        return input_
    assert input_.shape[dim] % world == 0, (f"input dimension {dim} ({input_.shape[dim]}) is not divisible by tensor "
                                            f"parallel world size ({world})")
    chunk = input_.shape[dim] // world
    return torch.narrow(input_, dim, rank * chunk, chunk)


class _GatherTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, dim):
        ctx.dim = dim
        return _gather_tokens(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _drop_tokens(grad_output, ctx.dim).contiguous(), None


class _DropTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, dim):
        ctx.dim = dim
        return _drop_tokens(input_, dim)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_tokens(grad_output, ctx.dim), None


def gather_tokens(input_, dim=0):
    return input_ if _tp()[0] == 1 else _GatherTokens.apply(input_, dim)


def drop_tokens(input_, dim=0):
    return input_ if _tp()[0] == 1 else _DropTokens.apply(input_, dim)
