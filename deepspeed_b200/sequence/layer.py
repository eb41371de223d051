"""DeepSpeed-Ulysses sequence parallelism: all-to-all between sequence-sharded and head-sharded layouts.

Reference: ``deepspeed/sequence/layer.py`` (``single_all_to_all`` :241-274, ``_SeqAllToAll`` :276,
``DistributedAttention`` :297-460, uneven heads :131-238, ``_DimZeroAllToAll`` :289). Activations are sharded on
the sequence dimension everywhere except inside attention, where q/k/v are re-sharded to heads so every rank runs
full-sequence attention on ``h/P`` heads.

B200 path: when symmetric memory is available the re-shard is ONE kernel (``csrc/symm.cu`` all-to-all, peers pull
their chunk over NVLink and the ``[P, s/P, ...] -> [s, ...]`` interleave is folded into the store addressing), so the
permute/contiguous passes of the reference (``layer.py:25-62``) disappear. Otherwise NCCL ``all_to_all_single``.
"""
import torch

from .. import comm as dist


def _world(group):
    return dist.get_world_size(group) if dist.is_initialized() else 1


def _a2a_scatter_heads_gather_seq(x, group, symm=None):
    """[B, s_local, H, D] (sequence sharded) -> [B, s_local*P, H/P, D] (head sharded)."""
    P = _world(group)
    if P == 1:
        return x
    B, s, H, D = x.shape
    if H % P != 0:
        return _uneven_scatter_heads_gather_seq(x, group)
    hp = H // P
    # chunk p (destination rank p) = heads [p*hp, (p+1)*hp): layout [P, B, s, hp, D]
    send = x.reshape(B, s, P, hp, D).permute(2, 0, 1, 3, 4).contiguous()
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv.view(-1), send.view(-1), group=group)
    # recv[p] = rank p's sequence chunk for my heads -> concatenate along sequence
    return recv.permute(1, 0, 2, 3, 4).reshape(B, P * s, hp, D)


def _a2a_scatter_seq_gather_heads(x, group):
    """[B, S, H/P, D] (head sharded) -> [B, S/P, H, D] (sequence sharded)."""
    P = _world(group)
    if P == 1:
        return x
    B, S, hp, D = x.shape
    s = S // P
    send = x.reshape(B, P, s, hp, D).permute(1, 0, 2, 3, 4).contiguous()  # [P, B, s, hp, D]
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv.view(-1), send.view(-1), group=group)
    # recv[p] = heads of rank p for my sequence chunk -> concatenate along heads
    return recv.permute(1, 2, 0, 3, 4).reshape(B, s, P * hp, D)


def head_shard_sizes(num_heads, world):
    """Heads per rank when ``num_heads`` is not divisible by the sequence-parallel size: the first ``num_heads % world``
    ranks take one extra head (reference ``get_shard_size_list``)."""
    base, extra = divmod(num_heads, world)
    return [base + (1 if r < extra else 0) for r in range(world)]


def _uneven_scatter_heads_gather_seq(x, group):
    """[B, s, H, D] -> [B, P*s, h_me, D] with unevenly split heads (split sizes on a heads-major layout)."""
    P, me = _world(group), dist.get_rank(group)
    B, s, H, D = x.shape
    hs = head_shard_sizes(H, P)
    send = x.permute(2, 0, 1, 3).contiguous()  # [H, B, s, D]
    recv = torch.empty((P * hs[me], B, s, D), dtype=x.dtype, device=x.device)
    dist.all_to_all_single(recv, send, output_split_sizes=[hs[me]] * P, input_split_sizes=hs, group=group)
    return recv.view(P, hs[me], B, s, D).permute(2, 0, 3, 1, 4).reshape(B, P * s, hs[me], D)


def _uneven_scatter_seq_gather_heads(x, group, num_heads):
    """[B, P*s, h_me, D] -> [B, s, H, D]: inverse of the above."""
    P, me = _world(group), dist.get_rank(group)
    B, S, h_me, D = x.shape
    hs = head_shard_sizes(num_heads, P)
    s = S // P
    send = x.reshape(B, P, s, h_me, D).permute(1, 3, 0, 2, 4).contiguous().view(P * h_me, B, s, D)
    recv = torch.empty((num_heads, B, s, D), dtype=x.dtype, device=x.device)
    dist.all_to_all_single(recv, send, output_split_sizes=hs, input_split_sizes=[h_me] * P, group=group)
    return recv.permute(1, 2, 0, 3).contiguous()


def uneven_heads_all2all(input, scatter_idx, gather_idx, batch_dim_idx, group):
    """Reference ``sequence/layer.py:131``: the Ulysses all-to-all when the head count is not divisible by the group
    size. ``scatter_idx >= 2`` scatters heads and gathers the sequence; otherwise the reverse, for which the full head
    count is recovered from the group's shard sizes. Layouts ``[b, s, h, d]`` (``batch_dim_idx`` 0) or ``[s, b, h, d]``."""
    x = input if batch_dim_idx == 0 else input.transpose(0, 1)
    if scatter_idx >= 2:
        out = _uneven_scatter_heads_gather_seq(x, group)
    else:
        P = _world(group)
        mine = torch.tensor([x.shape[2]], dtype=torch.long, device=x.device)
        all_h = [torch.zeros_like(mine) for _ in range(P)]
        dist.all_gather(all_h, mine, group=group)
        out = _uneven_scatter_seq_gather_heads(x, group, int(sum(int(t) for t in all_h)))
    return out if batch_dim_idx == 0 else out.transpose(0, 1).contiguous()


def pre_all2all_fun(permute_idx, inp_shape, input):
    """Reshape (and optionally permute) before an all-to-all (reference ``sequence/layer.py:80``)."""
    t = input.reshape(inp_shape).contiguous()
    return t.permute(permute_idx).contiguous() if permute_idx is not None else t


def post_all2all(permute_idx, res_shape):
    """The matching post-processing closure (reference ``sequence/layer.py:65``)."""

    def post_func(input):
        if permute_idx is not None:
            input = input.permute(permute_idx).contiguous()
        return input.reshape(res_shape).contiguous()

    return post_func


def apply_rotary_pos_emb(t, freqs_cos, freqs_sin):
    """Rotary embedding in the half-rotation convention on the first ``freqs.shape[-1]`` channels of ``t`` (tables
    broadcast against ``t``); reference ``sequence/layer.py:99``."""
    rot = freqs_cos.shape[-1]
    x, rest = t[..., :rot], t[..., rot:]
    x1, x2 = x[..., :rot // 2], x[..., rot // 2:]
    x = x * freqs_cos + torch.cat((-x2, x1), dim=-1) * freqs_sin
    return x if rest.shape[-1] == 0 else torch.cat((x, rest), dim=-1)


class _SeqAllToAll(torch.autograd.Function):
    """Differentiable layout switch; backward is the inverse all-to-all."""

    @staticmethod
    def forward(ctx, group, x, to_heads, num_heads=None):
        ctx.group, ctx.to_heads = group, to_heads
        P = _world(group)
        # uneven head split: the full head count travels with the call (it cannot be read off a head-sharded tensor)
        ctx.num_heads = x.shape[2] if to_heads else num_heads
        ctx.uneven = P > 1 and ctx.num_heads is not None and ctx.num_heads % P != 0
        if to_heads:
            return _a2a_scatter_heads_gather_seq(x, group)
        if ctx.uneven:
            return _uneven_scatter_seq_gather_heads(x, group, num_heads)
        return _a2a_scatter_seq_gather_heads(x, group)

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        if ctx.to_heads:
            out = _uneven_scatter_seq_gather_heads(g, ctx.group, ctx.num_heads) if ctx.uneven else \
                _a2a_scatter_seq_gather_heads(g, ctx.group)
        else:
            out = _a2a_scatter_heads_gather_seq(g, ctx.group)
        return None, out, None, None


def seq_to_heads(x, group):
    return _SeqAllToAll.apply(group, x, True)


def heads_to_seq(x, group, num_heads=None):
    """``num_heads``: total head count, needed only when it is not divisible by the group size."""
    return _SeqAllToAll.apply(group, x, False, num_heads)


def single_all_to_all(input, scatter_idx, gather_idx, batch_dim_idx, group, async_op=False, handle=None, type=None):
    """Reference-compatible helper for ``[s, b, h, d]`` / ``[b, s, h, d]`` tensors (scatter one dim, gather another)."""
    P = _world(group)
    if P == 1:
        return input
    if batch_dim_idx == 0:  # [b, s, h, d]
        if scatter_idx == 2:
            return _a2a_scatter_heads_gather_seq(input, group)
        return _a2a_scatter_seq_gather_heads(input, group)
    x = input.transpose(0, 1)  # [s, b, h, d] -> [b, s, h, d]
    out = _a2a_scatter_heads_gather_seq(x, group) if scatter_idx == 2 else _a2a_scatter_seq_gather_heads(x, group)
    return out.transpose(0, 1).contiguous()


class _DimZeroAllToAll(torch.autograd.Function):
    """Equal-split all-to-all over dim 0 (used by the HF/ALST flavour)."""

    @staticmethod
    def forward(ctx, group, x):
        ctx.group = group
        x = x.contiguous()
        out = torch.empty_like(x)
        dist.all_to_all_single(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        return None, _DimZeroAllToAll.apply(ctx.group, g)


class DistributedAttention(torch.nn.Module):
    """Wrap any local attention ``f(q, k, v, *args)`` so it runs under Ulysses sequence parallelism.

    Arguments follow the reference: ``local_attention``, ``sequence_process_group``, ``scatter_idx=2``,
    ``gather_idx=0`` for ``[s, b, h, d]`` inputs (``batch_dim_idx=1``) or ``[b, s, h, d]`` (``batch_dim_idx=0``).
    K/V with fewer heads than the SP degree are replicated (GQA/MQA, reference ``ulysses_sp.py:181-184``).
    """

    def __init__(self, local_attention, sequence_process_group, scatter_idx=2, gather_idx=0, sp_stream=None):
        super().__init__()
        self.local_attn = local_attention
        self.spg = sequence_process_group
        self.scatter_idx, self.gather_idx = scatter_idx, gather_idx
        self.sp_stream = sp_stream
        self.world = _world(sequence_process_group)

    def layer_sync(self, layer):
        """Wait for a projection that ran on the side stream (``sp_stream`` overlap) before its output is used."""
        ev = getattr(layer, "done_event", None)
        if self.sp_stream is not None and ev is not None and torch.cuda.is_available():
            torch.cuda.current_stream().wait_event(ev)

    def _prep_kv(self, t, head_dim_idx):
        H = t.shape[head_dim_idx]
        if H % self.world != 0 and self.world % H == 0:  # fewer kv heads than ranks: replicate (GQA / MQA)
            t = t.repeat_interleave(self.world // H, dim=head_dim_idx)
        return t  # otherwise the heads are split unevenly, the same way for q, k and v

    def forward(self, query, key, value, batch_dim_idx=1, *args, **kwargs):
        if self.world == 1:
            return self.local_attn(query, key, value, *args, **kwargs)
        bsz_first = batch_dim_idx == 0
        to_bshd = (lambda t: t) if bsz_first else (lambda t: t.transpose(0, 1))
        q = seq_to_heads(to_bshd(query).contiguous(), self.spg)
        k = seq_to_heads(self._prep_kv(to_bshd(key), 2).contiguous(), self.spg)
        v = seq_to_heads(self._prep_kv(to_bshd(value), 2).contiguous(), self.spg)
        if not bsz_first:
            q, k, v = q.transpose(0, 1), k.transpose(0, 1), v.transpose(0, 1)
        ctx = self.local_attn(q, k, v, *args, **kwargs)
        ctx_b = ctx if bsz_first else ctx.transpose(0, 1)
        if ctx_b.dim() == 3:  # [b, s, h*d] -> split heads
            hp = q.shape[2] if bsz_first else q.shape[2]
            ctx_b = ctx_b.reshape(ctx_b.shape[0], ctx_b.shape[1], hp, -1)
        out = heads_to_seq(ctx_b.contiguous(), self.spg, num_heads=to_bshd(query).shape[2])
        return out if bsz_first else out.transpose(0, 1).contiguous()


class UlyssesAttention:
    """Callable used by :class:`deepspeed_b200.models.llama.LlamaModel` (``sp_attention``): q/k/v arrive sequence
    sharded ``[B, S/P, H, D]`` (RoPE already applied with global positions)."""

    def __init__(self, group, causal=True, attn_impl="auto"):
        self.group, self.causal, self.attn_impl = group, causal, attn_impl
        self.world = _world(group)

    def __call__(self, q, k, v):
        from ..models.llama import attention
        P = self.world
        if k.shape[2] % P != 0:
            rep = P // k.shape[2]
            k, v = k.repeat_interleave(rep, dim=2), v.repeat_interleave(rep, dim=2)
        qh, kh, vh = seq_to_heads(q.contiguous(), self.group), seq_to_heads(k.contiguous(), self.group), \
            seq_to_heads(v.contiguous(), self.group)
        o = attention(qh, kh, vh, causal=self.causal, impl=self.attn_impl)
        return heads_to_seq(o.contiguous(), self.group)
