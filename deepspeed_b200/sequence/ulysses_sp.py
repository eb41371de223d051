"""Ulysses/ALST (Arctic Long Sequence Training) integration for Hugging Face style models + sequence tiling.

Reference: ``runtime/sequence_parallel/ulysses_sp.py`` -- ``UlyssesSPAttentionHF`` :50-561 (registers itself as an
``ALL_ATTENTION_FUNCTIONS`` entry, GQA/MQA kv replication :131-184, position-ids based causal masking),
``UlyssesSPDataLoaderAdapter`` :564-717, ``SequenceTiledCompute`` :720-941, ``TiledMLP`` :943-1063,
``TiledFusedLogitsLoss`` :1065-1216, ``UlyssesSPFwdLossBwdWithLogits`` :1418.
"""
import math
from collections import defaultdict

import torch
import torch.nn.functional as F

from .. import comm as dist
from ..utils import groups
from .layer import heads_to_seq, seq_to_heads


class UlyssesSPAttentionHF(torch.nn.Module):
    """Wraps an HF attention function ``attn(module, q, k, v, attention_mask, **kw)`` with q/k/v ``[b, h, s/P, d]``."""

    def __init__(self, attn, local_seq_length=None, global_seq_length=None, batch_size=None, attn_head_count=None,
                 attn_head_size=None, kv_head_count=None, num_hidden_layers=None, process_group=None,
                 seq_length_is_variable=True):
        super().__init__()
        self.attn = attn
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.sp_rank = dist.get_rank(process_group) if dist.is_initialized() else 0
        self.attn_head_count, self.kv_head_count = attn_head_count, kv_head_count
        self.attn_head_size = attn_head_size
        self.seq_length_is_variable = seq_length_is_variable
        if attn_head_count is not None and self.world_size > 1:
            if attn_head_count % self.world_size != 0:
                raise ValueError(f"Attention head count {attn_head_count} is not divisible by SP size {self.world_size}")
            if kv_head_count is not None and not (kv_head_count % self.world_size == 0
                                                   or self.world_size % kv_head_count == 0):
                raise ValueError(f"kv head count {kv_head_count} and SP size {self.world_size} must divide one another")

    def forward(self, module, query, key, value, attention_mask=None, *args, **kwargs):
        if self.world_size == 1:
            return self.attn(module, query, key, value, attention_mask, *args, **kwargs)
        # HF layout [b, h, s, d] -> [b, s, h, d]
        q, k, v = (t.transpose(1, 2).contiguous() for t in (query, key, value))
        P = self.world_size
        if k.shape[2] % P != 0:
            rep = P // k.shape[2]
            k, v = k.repeat_interleave(rep, dim=2), v.repeat_interleave(rep, dim=2)
        q, k, v = seq_to_heads(q, self.process_group), seq_to_heads(k, self.process_group), \
            seq_to_heads(v, self.process_group)
        if "position_ids" in kwargs and kwargs["position_ids"] is not None:
            pos = kwargs["position_ids"]
            gathered = [torch.empty_like(pos) for _ in range(P)]
            dist.all_gather(gathered, pos.contiguous(), group=self.process_group)
            kwargs["position_ids"] = torch.cat(gathered, dim=1)
        out = self.attn(module, q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), None, *args, **kwargs)
        attn_out, rest = (out[0], out[1:]) if isinstance(out, tuple) else (out, ())
        # HF attention functions return [b, s, h, d]
        if attn_out.dim() == 3:
            attn_out = attn_out.reshape(attn_out.shape[0], attn_out.shape[1], q.shape[2], -1)
        attn_out = heads_to_seq(attn_out.contiguous(), self.process_group)
        return (attn_out, ) + tuple(rest) if rest else (attn_out, None)

    @classmethod
    def register_with_transformers(cls, model_name_or_path, core_attn_implementation, sequence_parallel_size,
                                   micro_batch_size=None, seq_length=None, seq_length_is_variable=True):
        """Install the wrapper into ``transformers.ALL_ATTENTION_FUNCTIONS`` under the name ``ulysses``."""
        if sequence_parallel_size == 1:
            return None
        from transformers import AutoConfig
        from transformers.modeling_utils import ALL_ATTENTION_FUNCTIONS
        cfg = model_name_or_path if not isinstance(model_name_or_path, str) else AutoConfig.from_pretrained(
            model_name_or_path)
        core = ALL_ATTENTION_FUNCTIONS[core_attn_implementation]
        groups.initialize(sp_size=sequence_parallel_size)
        obj = cls(attn=core, attn_head_count=cfg.num_attention_heads,
                  attn_head_size=getattr(cfg, "head_dim", cfg.hidden_size // cfg.num_attention_heads),
                  kv_head_count=getattr(cfg, "num_key_value_heads", cfg.num_attention_heads),
                  num_hidden_layers=cfg.num_hidden_layers, process_group=groups._get_sequence_parallel_group(),
                  seq_length_is_variable=seq_length_is_variable)

        def wrapper(module, query, key, value, attention_mask, *args, **kwargs):
            return obj(module, query, key, value, attention_mask, *args, **kwargs)

        ALL_ATTENTION_FUNCTIONS["ulysses"] = wrapper
        return groups


class UlyssesSPDataLoaderAdapter:
    """Each DP sample is processed by all SP ranks: every rank's batch is broadcast in turn and sharded on the
    sequence dimension (``input_ids``, ``position_ids``, ``shift_labels``, ...)."""

    def __init__(self, dl, sp_rank, sp_group, sp_world_size, device):
        self.dl, self.sp_rank, self.sp_group, self.sp_world_size, self.device = dl, sp_rank, sp_group, sp_world_size, device
        self.iter = iter(dl)
        self.micro_batches = []

    def __len__(self):
        return len(self.dl) * self.sp_world_size

    def __iter__(self):
        return self

    def __next__(self):
        if not self.micro_batches:
            self.refill()
        return self.micro_batches.pop(0)

    def refill(self):
        batch = next(self.iter)
        for k in list(batch.keys()):
            if torch.is_tensor(batch[k]):
                batch[k] = batch[k].to(self.device)
        if "labels" in batch and "shift_labels" not in batch:
            lab = F.pad(batch["labels"], (0, 1), value=-100)
            batch["shift_labels"] = lab[..., 1:].contiguous()
            del batch["labels"]
        if "position_ids" not in batch and "input_ids" in batch:
            b, s = batch["input_ids"].shape
            batch["position_ids"] = torch.arange(s, device=self.device).unsqueeze(0).expand(b, s).contiguous()
        batch.pop("attention_mask", None)  # causal masking is position-id driven under SP
        micro = defaultdict(dict)
        for k, t in batch.items():
            if not torch.is_tensor(t):
                for r in range(self.sp_world_size):
                    micro[r][k] = t
                continue
            gathered = [torch.empty_like(t) for _ in range(self.sp_world_size)] if self.sp_world_size > 1 else [t]
            if self.sp_world_size > 1:
                shape = torch.tensor(list(t.shape), device=self.device)
                dist.all_gather(gathered, t.contiguous(), group=self.sp_group)
            for r in range(self.sp_world_size):
                full = gathered[r]
                if full.dim() >= 2 and full.shape[1] % self.sp_world_size == 0:
                    chunk = full.shape[1] // self.sp_world_size
                    micro[r][k] = full[:, self.sp_rank * chunk:(self.sp_rank + 1) * chunk].contiguous()
                else:
                    micro[r][k] = full
        self.micro_batches = [micro[r] for r in range(self.sp_world_size)]


def sequence_tiled_compute(fn, seqlen, shards, kwargs_to_shard, kwargs_to_pass, grad_requiring_tensor_key,
                           compute_params=None, output_unshard_dimension=1, output_reduction="mean"):
    """Functional front of :class:`SequenceTiledCompute`."""
    return SequenceTiledCompute.apply(fn, seqlen, shards, list(kwargs_to_shard.keys()), list(kwargs_to_pass.keys()),
                                      grad_requiring_tensor_key, compute_params, output_unshard_dimension,
                                      output_reduction, *kwargs_to_shard.values(), *kwargs_to_pass.values())


class SequenceTiledCompute(torch.autograd.Function):
    """Run ``fn(**kwargs)`` tile by tile over the sequence dimension so only one tile's activations are alive; the
    backward recomputes each tile (memory O(tile) instead of O(sequence))."""

    @staticmethod
    def forward(ctx, fn, seqlen, shards, keys_to_shard, keys_to_pass, grad_requiring_tensor_key, compute_params,
                output_unshard_dimension, output_reduction, *args):
        ctx.fn, ctx.seqlen, ctx.shards = fn, seqlen, shards
        ctx.keys_to_shard, ctx.keys_to_pass = keys_to_shard, keys_to_pass
        ctx.grad_key, ctx.compute_params = grad_requiring_tensor_key, compute_params
        ctx.dim, ctx.reduction = output_unshard_dimension, output_reduction
        ns = len(keys_to_shard)
        ctx.args = args
        to_shard = dict(zip(keys_to_shard, args[:ns]))
        to_pass = dict(zip(keys_to_pass, args[ns:]))
        outs = []
        with torch.no_grad():
            sharded = {k: list(torch.tensor_split(v, shards, dim=1)) for k, v in to_shard.items()}
            ctx.tile_sizes = [t.shape[1] for t in sharded[keys_to_shard[0]]]
            for i in range(shards):
                outs.append(fn(**{k: v[i] for k, v in sharded.items()}, **to_pass))
        if output_reduction == "mean":
            w = torch.tensor(ctx.tile_sizes, dtype=outs[0].dtype, device=outs[0].device) / float(sum(ctx.tile_sizes))
            return (torch.stack([o.reshape(()) for o in outs]) * w).sum()
        if output_reduction == "sum":
            return torch.stack([o.reshape(()) for o in outs]).sum()
        return torch.cat(outs, dim=output_unshard_dimension)

    @staticmethod
    def backward(ctx, *grads):
        g = grads[0]
        ns = len(ctx.keys_to_shard)
        args = ctx.args
        to_shard = dict(zip(ctx.keys_to_shard, args[:ns]))
        to_pass = dict(zip(ctx.keys_to_pass, args[ns:]))
        x = to_shard[ctx.grad_key]
        x_grad = torch.zeros_like(x)
        total = float(sum(ctx.tile_sizes))
        off = 0
        sharded = {k: list(torch.tensor_split(v, ctx.shards, dim=1)) for k, v in to_shard.items()}
        g_tiles = list(torch.tensor_split(g, ctx.shards, dim=ctx.dim)) if ctx.reduction not in ("mean", "sum") else None
        for i in range(ctx.shards):
            kw = {k: v[i] for k, v in sharded.items()}
            xi = kw[ctx.grad_key].detach().requires_grad_(True)
            kw[ctx.grad_key] = xi
            with torch.enable_grad():
                out = ctx.fn(**kw, **to_pass)
            if ctx.reduction == "mean":
                out.backward(g * (ctx.tile_sizes[i] / total))
            elif ctx.reduction == "sum":
                out.backward(g)
            else:
                out.backward(g_tiles[i])
            x_grad.narrow(1, off, ctx.tile_sizes[i]).copy_(xi.grad)
            off += ctx.tile_sizes[i]
        out_grads = [None] * 9
        for k in ctx.keys_to_shard:
            out_grads.append(x_grad if k == ctx.grad_key else None)
        out_grads += [None] * len(ctx.keys_to_pass)
        return tuple(out_grads)


class TiledMLP(torch.autograd.Function):
    """``fn(self, x)`` over sequence tiles of ``x`` ``[b, s, h]`` with per-tile recompute in backward."""

    @staticmethod
    def forward(ctx, fn, self_mod, x, shards, compute_params=None):
        ctx.fn, ctx.self_mod, ctx.shards, ctx.compute_params = fn, self_mod, shards, compute_params
        ctx.save_for_backward(x)
        with torch.no_grad():
            outs = [fn(self_mod, t) for t in torch.tensor_split(x, shards, dim=1)]
        return torch.cat(outs, dim=1)

    @staticmethod
    def backward(ctx, g):
        (x, ) = ctx.saved_tensors
        xg = torch.zeros_like(x)
        off = 0
        for xt, gt in zip(torch.tensor_split(x, ctx.shards, dim=1), torch.tensor_split(g, ctx.shards, dim=1)):
            xi = xt.detach().requires_grad_(True)
            with torch.enable_grad():
                out = ctx.fn(ctx.self_mod, xi)
            torch.autograd.backward(out, gt)
            xg.narrow(1, off, xt.shape[1]).copy_(xi.grad)
            off += xt.shape[1]
        return None, None, xg, None, None


def tiled_mlp_forward(mlp_module, x, shards=None, min_tile_tokens=2048):
    """Convenience: tile ``mlp_module.forward`` over the sequence (auto number of shards)."""
    if shards is None:
        shards = max(1, math.ceil(x.shape[1] / max(min_tile_tokens, x.shape[-1])))
    params = list(mlp_module.parameters())
    return TiledMLP.apply(lambda m, t: type(m).forward(m, t), mlp_module, x, shards, params)


class TiledFusedLogitsLoss(torch.autograd.Function):
    """``loss_fn(self, x_tile, y_tile)`` over token tiles without ever materialising the full logits; gradients of
    ``x`` are produced during the forward (reference :1065-1216)."""

    @staticmethod
    def forward(ctx, fn, self_mod, x, y, mask, shards, compute_params, output_reduction="mean"):
        assert output_reduction in ("mean", "sum")
        b, s, h = x.shape
        xs, ys = x.reshape(-1, h), y.reshape(-1)
        xg = torch.zeros_like(xs)
        n_valid = (ys != -100).sum().clamp(min=1) if output_reduction == "mean" else None
        total = torch.zeros((), dtype=torch.float32, device=x.device)
        off = 0
        for xt, yt in zip(torch.tensor_split(xs, shards, dim=0), torch.tensor_split(ys, shards, dim=0)):
            xi = xt.detach().requires_grad_(True)
            with torch.enable_grad():
                loss_sum = fn(self_mod, xi, yt)  # SUM over the tile's valid tokens
                scaled = loss_sum / n_valid if n_valid is not None else loss_sum
            scaled.backward()
            xg[off:off + xt.shape[0]].copy_(xi.grad)
            total += scaled.detach().float()
            off += xt.shape[0]
        ctx.save_for_backward(xg.view(b, s, h))
        return total

    @staticmethod
    def backward(ctx, g):
        (xg, ) = ctx.saved_tensors
        return None, None, xg * g.to(xg.dtype), None, None, None, None, None


class UlyssesSPFwdLossBwdWithLogits:
    """Helper for models that return logits: computes the SP-aware mean loss over all ranks' valid tokens."""

    def __init__(self, model, model_unwrapped, device, num_loss_logit_shards="auto"):
        self.model, self.model_unwrapped, self.device = model, model_unwrapped, device
        self.num_loss_logit_shards = num_loss_logit_shards
        self.sp_group = groups._get_sequence_parallel_group()
        self.sp_world_size = groups._get_sequence_parallel_world_size()

    def sp_fwd_loss_bwd(self, batch):
        shift_labels = batch.pop("shift_labels")
        out = self.model(**batch, use_cache=False)
        logits = out.logits if hasattr(out, "logits") else out
        loss_sum = F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), shift_labels.reshape(-1),
                                   ignore_index=-100, reduction="sum")
        good = (shift_labels != -100).sum().float()
        stats = torch.stack([loss_sum.detach(), good])
        if self.sp_world_size > 1:
            dist.all_reduce(stats, group=self.sp_group)
        loss = loss_sum / stats[1].clamp(min=1) * self.sp_world_size  # grads are averaged over sp by ZeRO
        self.model.backward(loss)
        return stats[0] / stats[1].clamp(min=1)


class AutogradComputeMLP(torch.autograd.Function):
    """``AutogradComputeMLP.apply(fn, module, x)``: run ``fn(module, x)`` without keeping its activations and recompute
    them in backward - the un-tiled skeleton of :class:`TiledMLP` (reference
    ``runtime/sequence_parallel/ulysses_sp.py:1218``)."""

    @staticmethod
    def forward(ctx, fn, self, x):
        ctx.fn, ctx.self = fn, self
        ctx.save_for_backward(x)
        with torch.no_grad():
            return fn(self, x)

    @staticmethod
    def backward(ctx, *grads):
        (x, ) = ctx.saved_tensors
        x1 = x.detach().requires_grad_(x.requires_grad)
        with torch.enable_grad():
            out = ctx.fn(ctx.self, x1)
        torch.autograd.backward(out, grads[0])
        return None, None, x1.grad


class TiledLoss(torch.autograd.Function):
    """Loss over existing logits computed shard by shard along the sequence, so that the fp32 upcast / softmax
    temporaries exist for one shard at a time (reference ``ulysses_sp.py:1269``; :class:`TiledFusedLogitsLoss` goes
    further and never materialises the logits).

    ``loss_fn(logits=, labels=None, vocab_size=, shift_labels=)`` returns the mean over the labels that are not
    ``-100``. The shards' means are combined weighted by their number of such labels (a shard made only of ``-100`` is
    skipped), and backward weights each shard's gradient the same way, so the result and its gradient equal the
    un-tiled call whatever the distribution of ignored labels. (The reference divides the accumulated gradient by the
    shard count instead, which is exact only when every shard holds the same number of valid labels.)"""

    @staticmethod
    def _steps(seqlen, shards):
        step = -(-seqlen // shards)
        return [(lo, min(seqlen, lo + step)) for lo in range(0, seqlen, step)]

    @staticmethod
    def forward(ctx, loss_fn, logits, vocab_size, shift_labels, shards):
        ctx.save_for_backward(logits, shift_labels)
        ctx.loss_fn, ctx.vocab_size, ctx.shards = loss_fn, vocab_size, shards
        total = None
        counts = []
        with torch.no_grad():
            for lo, hi in TiledLoss._steps(shift_labels.shape[1], shards):
                lab = shift_labels[:, lo:hi]
                n = int((lab != -100).sum())
                counts.append(n)
                if n == 0:
                    continue
                part = loss_fn(logits=logits[:, lo:hi, :], labels=None, vocab_size=vocab_size, shift_labels=lab) * n
                total = part if total is None else total + part
        ctx.counts = counts
        n_all = sum(counts)
        if n_all == 0:
            return (logits.sum() * 0.0).float()
        return total / n_all

    @staticmethod
    def backward(ctx, *grads):
        logits, shift_labels = ctx.saved_tensors
        g = grads[0]
        n_all = sum(ctx.counts)
        logits_grad = torch.zeros_like(logits)
        if n_all > 0:
            for (lo, hi), n in zip(TiledLoss._steps(shift_labels.shape[1], ctx.shards), ctx.counts):
                if n == 0:
                    continue
                shard = logits[:, lo:hi, :].detach().requires_grad_(True)
                with torch.enable_grad():
                    loss = ctx.loss_fn(logits=shard, labels=None, vocab_size=ctx.vocab_size,
                                       shift_labels=shift_labels[:, lo:hi])
                (gs, ) = torch.autograd.grad(loss, shard, g * (n / n_all))
                logits_grad[:, lo:hi, :] = gs
        return None, logits_grad, None, None, None
