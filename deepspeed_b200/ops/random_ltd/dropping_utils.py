"""Random layerwise token dropping ops (reference ``ops/random_ltd/dropping_utils.py``; kernels
``csrc/random_ltd/*.cu``): sample a sorted subset of token positions per (layer, sample), gather those tokens for the
layer, scatter the layer output back. On CUDA the row movement and the index sort run on ``csrc/token_ops.cu``."""
import torch

from .. import kernels as K


def _on_gpu(t):
    return t.is_cuda and K.available()


def _sample_sorted(reserved_length, seq_length, batch_size, layers, device):
    prob = torch.ones((layers * batch_size, seq_length), device=device)
    idx = torch.multinomial(prob, reserved_length).to(torch.int32)
    if _on_gpu(idx) and reserved_length <= 8192:
        K.sort_rows_i32_(idx)
    else:
        idx = torch.sort(idx, dim=-1).values
    return idx.reshape(layers, batch_size, reserved_length)


def gpt_sample_tokens(reserved_length, seq_length, batch_size, layers=1, device="cpu", attn_mask=None):
    """-> (sampled_indices [layers, batch, reserved], causal mask cropped to [b, 1, reserved, reserved])."""
    idx = _sample_sorted(reserved_length, seq_length, batch_size, layers, device)
    new_mask = attn_mask[:, :, :reserved_length, :reserved_length] if attn_mask is not None else None
    return idx, new_mask


def bert_sample_tokens(reserved_length, seq_length, batch_size, layers=1, device="cpu", attn_mask=None):
    """-> (sampled_indices, per-layer masks [b, 1, reserved, reserved] restricted to the kept tokens)."""
    assert attn_mask is not None
    idx = _sample_sorted(reserved_length, seq_length, batch_size, layers, device)
    li = idx.long()
    masks = []
    for layer in range(layers):
        rows = torch.gather(attn_mask, 2, li[layer][:, None, :, None].expand(-1, attn_mask.shape[1], -1,
                                                                            attn_mask.shape[3]))
        masks.append(torch.gather(rows, 3, li[layer][:, None, None, :].expand(-1, rows.shape[1], rows.shape[2], -1)))
    return idx, masks


def _flat_rows(indices, batch_first, seq, batch):
    """Row numbers into the activation viewed as [batch*seq, H] (batch_first) or [seq*batch, H]."""
    idx = indices.long()
    b = torch.arange(batch, device=idx.device)[:, None]
    return (b * seq + idx).reshape(-1) if batch_first else (idx * batch + b).t().reshape(-1)


def token_gather(activations, sorted_indices, batch_first):
    if batch_first:
        B, S, H = activations.shape
    else:
        S, B, H = activations.shape
    R = sorted_indices.shape[-1]
    rows = _flat_rows(sorted_indices, batch_first, S, B)
    src = activations.reshape(-1, H)
    if _on_gpu(src) and (H * src.element_size()) % 16 == 0 and src.is_contiguous():
        out = K.gather_rows(src, rows)
    else:
        out = src.index_select(0, rows)
    return out.view(B, R, H) if batch_first else out.view(R, B, H)


def token_scatter_(all_activations, layer_activations, sorted_indices, batch_first):
    if batch_first:
        B, S, H = all_activations.shape
    else:
        S, B, H = all_activations.shape
    rows = _flat_rows(sorted_indices, batch_first, S, B)
    dst = all_activations.view(-1, H)
    src = layer_activations.reshape(-1, H)
    if _on_gpu(dst) and (H * dst.element_size()) % 16 == 0 and src.is_contiguous():
        K.scatter_rows_(dst, rows, src)
    else:
        dst.index_copy_(0, rows, src)
    return all_activations


class GatherTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, activations, sorted_indices, batch_first):
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return activations, token_gather(activations, sorted_indices, batch_first)

    @staticmethod
    def backward(ctx, a_gradients, g_gradients):
        sorted_indices, = ctx.saved_tensors
        # the kept positions receive the layer's gradient instead of the bypass gradient (the scatter in forward
        # overwrote those rows, so their bypass gradient is zero by construction of ScatterTokens.backward)
        out = a_gradients.contiguous().clone()
        return token_scatter_(out, g_gradients.contiguous(), sorted_indices, ctx.batch_first), None, None


class ScatterTokens(torch.autograd.Function):

    @staticmethod
    def forward(ctx, all_activations, layer_activations, sorted_indices, batch_first):
        ctx.save_for_backward(sorted_indices)
        ctx.batch_first = batch_first
        return token_scatter_(all_activations.contiguous().clone(), layer_activations.contiguous(), sorted_indices,
                              batch_first)

    @staticmethod
    def backward(ctx, out_gradients):
        sorted_indices, = ctx.saved_tensors
        out_gradients = out_gradients.contiguous()
        return out_gradients, token_gather(out_gradients, sorted_indices, ctx.batch_first), None, None
