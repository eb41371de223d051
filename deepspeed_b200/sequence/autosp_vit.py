"""Sequence-parallel wrapper for ViT encoder attention (reference ``sequence/autosp_vit.py``).

ViT attention is bidirectional, so the wrapped attention must see every patch: patches arrive sharded along the
sequence, are all-gathered (uneven shards allowed), the wrapped module runs on the full sequence, and each rank keeps
its own slice of the output. Everything outside the attention (MLP, norms) stays sequence-sharded, which is where the
activation memory goes. A [CLS] token, when present, is replicated on every rank and never sharded.

The gather is differentiable here (backward: sum the gradient slices of all ranks and keep the local one), so the
wrapper can be trained through."""
import torch
import torch.nn as nn

from .. import comm as dist


class _GatherPatches(torch.autograd.Function):
    """[b, n_r, h] per rank -> [b, sum n_r, h] on every rank."""

    @staticmethod
    def forward(ctx, x, group, lens):
        world = len(lens)
        rank = dist.get_rank(group)
        width = max(lens)
        mine = x if x.shape[1] == width else torch.nn.functional.pad(x, (0, 0, 0, width - x.shape[1]))
        bufs = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(bufs, mine.contiguous(), group=group)
        ctx.group, ctx.lens, ctx.rank = group, lens, rank
        return torch.cat([bufs[r][:, :lens[r]] for r in range(world)], dim=1)

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().clone()
        dist.all_reduce(g, group=ctx.group)  # every rank used every patch
        start = sum(ctx.lens[:ctx.rank])
        return g[:, start:start + ctx.lens[ctx.rank]].contiguous(), None, None


class UlyssesSPViTAttention(nn.Module):

    def __init__(self, attn: nn.Module, process_group, has_cls_token: bool = True):
        super().__init__()
        self.attn = attn
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group)
        self.has_cls_token = has_cls_token

    def _shard_lengths(self, n_local, device):
        mine = torch.tensor([n_local], dtype=torch.long, device=device)
        out = [torch.zeros_like(mine) for _ in range(self.world_size)]
        dist.all_gather(out, mine, group=self.process_group)
        return [int(t) for t in out]

    def forward(self, hidden_states, **kwargs):
        cls_tok = hidden_states[:, :1] if self.has_cls_token else None
        patches = hidden_states[:, 1:] if self.has_cls_token else hidden_states
        lens = self._shard_lengths(patches.shape[1], patches.device)
        full = _GatherPatches.apply(patches, self.process_group, lens)
        if cls_tok is not None:
            full = torch.cat([cls_tok, full], dim=1)
        out = self.attn(full, **kwargs)
        extra = ()
        if isinstance(out, (tuple, list)):
            out, extra = out[0], tuple(out[1:])
        rank = dist.get_rank(self.process_group)
        first = (1 if self.has_cls_token else 0) + sum(lens[:rank])
        local = out[:, first:first + lens[rank]]
        if self.has_cls_token:
            local = torch.cat([out[:, :1], local], dim=1)
        local = local.contiguous()
        return (local, ) + extra if extra else local
