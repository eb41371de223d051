"""Vocab-parallel cross entropy over the sequence-parallel group (reference ``sequence/cross_entropy.py:11``)."""
import torch

from .. import comm as dist


class _VocabSequenceParallelCrossEntropy(torch.autograd.Function):

    @staticmethod
    def forward(ctx, vocab_seq_parallel_logits, target, sp_group):
        # logits: [S/P, B, V]; every rank needs the loss of the full sequence
        logits = vocab_seq_parallel_logits.float()
        logits_max = logits.max(dim=-1)[0]
        logits = logits - logits_max.unsqueeze(-1)
        sum_exp = logits.exp().sum(dim=-1)
        tgt = target.unsqueeze(-1)
        pred = logits.gather(-1, tgt).squeeze(-1)
        loss = torch.log(sum_exp) - pred
        ctx.save_for_backward(logits.exp() / sum_exp.unsqueeze(-1), target)
        world = dist.get_world_size(sp_group) if dist.is_initialized() else 1
        ctx.sp_group, ctx.world, ctx.rank = sp_group, world, (dist.get_rank(sp_group) if world > 1 else 0)
        ctx.local_s = loss.shape[0]
        if world > 1:
            gathered = [torch.empty_like(loss) for _ in range(world)]
            dist.all_gather(gathered, loss.contiguous(), group=sp_group)
            loss = torch.cat(gathered, dim=0)
        return loss

    @staticmethod
    def backward(ctx, grad_output):
        softmax, target = ctx.saved_tensors
        g = grad_output[ctx.rank * ctx.local_s:(ctx.rank + 1) * ctx.local_s]
        grad = softmax.clone()
        grad.scatter_add_(-1, target.unsqueeze(-1), -torch.ones_like(grad[..., :1]))
        grad.mul_(g.unsqueeze(-1))
        return grad, None, None


def vocab_sequence_parallel_cross_entropy(vocab_parallel_logits, target, sp_group):
    return _VocabSequenceParallelCrossEntropy.apply(vocab_parallel_logits, target, sp_group)
