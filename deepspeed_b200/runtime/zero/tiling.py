"""``TiledLinear`` (reference ``runtime/zero/tiling.py:32``): split a big linear into in/out tiles so ZeRO-3 only
gathers one tile's weights at a time; ``copy_params_from`` imports a dense layer."""
import torch
import torch.nn as nn

from ..pipe.module import partition_uniform


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    sizes = [partitions[i + 1] - partitions[i] for i in range(len(partitions) - 1)]
    chunks = torch.split(tensor, sizes, dim=tensor.dim() - 1)
    return tuple(c.contiguous() for c in chunks) if contiguous_split_chunks else chunks


class TiledLinear(nn.Module):

    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if in_splits < 1 or in_splits > in_features:
            raise RuntimeError("in splits must be in range [1, in_features].")
        if out_splits < 1 or out_splits > out_features:
            raise RuntimeError("out splits must be in range [1, out_features].")
        self.in_features, self.out_features, self.use_bias = in_features, out_features, bias
        self.in_splits, self.out_splits = in_splits, out_splits
        self.input_is_already_split, self.combine_out_splits = input_is_already_split, combine_out_splits
        self.in_parts = partition_uniform(in_features, in_splits)
        self.out_parts = partition_uniform(out_features, out_splits)
        self.linears = nn.ModuleList()
        for out_id in range(out_splits):
            self.linears.append(nn.ModuleList())
            local_out = self.out_parts[out_id + 1] - self.out_parts[out_id]
            for in_id in range(in_splits):
                local_bias = bias if in_id == in_splits - 1 else False
                local_in = self.in_parts[in_id + 1] - self.in_parts[in_id]
                self.linears[out_id].append(linear_cls(local_in, local_out, bias=local_bias, **kwargs))
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def forward(self, input_):
        if self.in_splits > 1 and not self.input_is_already_split:
            inputs = split_tensor_along_last_dim(input_, self.in_parts)
        elif self.in_splits > 1:
            inputs = input_
        else:
            inputs = [input_]
        outputs = [None] * self.out_splits
        for out_id in range(self.out_splits):
            for in_id in range(self.in_splits):
                local = self.linears[out_id][in_id](inputs[in_id])
                outputs[out_id] = local if outputs[out_id] is None else outputs[out_id] + local
        return torch.cat(outputs, dim=-1) if self.combine_out_splits else outputs

    @torch.no_grad()
    def copy_params_from(self, other):
        assert other.weight.shape == (self.out_features, self.in_features)
        for row in range(self.out_splits):
            rs, re = self.out_parts[row], self.out_parts[row + 1]
            for col in range(self.in_splits):
                cs, ce = self.in_parts[col], self.in_parts[col + 1]
                local = self.linears[row][col]
                local.weight.copy_(other.weight[rs:re, cs:ce])
                if local.bias is not None:
                    local.bias.copy_(other.bias[rs:re])


class TiledLinearReturnBias(TiledLinear):
    """Megatron-style linear returning ``(output, bias)``."""

    def forward(self, input_):
        out = super().forward(input_)
        return out, None
