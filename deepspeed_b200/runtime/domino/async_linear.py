"""Tensor-parallel linears whose all-reduces are issued asynchronously so that they overlap other work (reference
``runtime/domino/async_linear.py:18-140``)."""
import torch
import torch.nn.functional as F
from torch.nn.parameter import Parameter

from ... import comm as dist


class DominoAsyncColumnParallelLinearImpl(torch.autograd.Function):
    """Column-parallel ``y = x W^T``: identity forward; the backward's input-gradient all-reduce is launched async and
    its handle parked in ``handle_dic[h_id]`` - the consumer waits right before it needs the gradient."""

    @staticmethod
    def forward(ctx, inp, weight, bias, handle_dic, h_id, group, world=None):
        ctx.save_for_backward(inp, weight)
        ctx.use_bias = bias is not None
        ctx.handle_dic, ctx.h_id, ctx.group = handle_dic, h_id, group
        ctx.world = world if world is not None else (dist.get_world_size(group) if dist.is_initialized() else 1)
        return F.linear(inp, weight, bias)

    @staticmethod
    def backward(ctx, grad_output):
        inp, weight = ctx.saved_tensors
        grad_input = grad_output.matmul(weight)
        if ctx.world > 1:
            ctx.handle_dic[ctx.h_id] = dist.all_reduce(grad_input, group=ctx.group, async_op=True)
        go = grad_output.reshape(-1, grad_output.shape[-1])
        grad_weight = go.t().matmul(inp.reshape(-1, inp.shape[-1]))
        grad_bias = go.sum(0) if ctx.use_bias else None
        return grad_input, grad_weight, grad_bias, None, None, None, None


class DominoAsyncColumnParallelLinear(torch.nn.Module):

    def __init__(self, input_size, output_size, _tp_group, config=None, init_method=None, bias=True,
                 tp_world_size=None, dtype=torch.float32, device=None):
        super().__init__()
        self.tp_group = _tp_group
        world = tp_world_size or (dist.get_world_size(_tp_group) if dist.is_initialized() else 1)
        self.tp_world = world
        self.input_size, self.output_size = input_size, output_size
        self.output_size_per_partition = output_size // world
        self.weight = Parameter(torch.empty(self.output_size_per_partition, input_size, dtype=dtype, device=device))
        (init_method or torch.nn.init.xavier_normal_)(self.weight)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size_per_partition, dtype=dtype, device=device))
        else:
            self.register_parameter("bias", None)

    def forward(self, input_, handle_dic, h_id):
        out = DominoAsyncColumnParallelLinearImpl.apply(input_, self.weight, self.bias, handle_dic, h_id,
                                                        self.tp_group, self.tp_world)
        return out, None


class RowParallelLinearNoComm(torch.nn.Module):
    """Row-parallel linear that leaves the all-reduce of its partial output to the caller (who overlaps it)."""

    def __init__(self, input_size, output_size, config=None, init_method=None, bias=True, stride=1,
                 skip_bias_add=False, tp_world_size=1, dtype=torch.float32, device=None):
        super().__init__()
        self.input_size, self.output_size = input_size, output_size
        self.input_size_per_partition = input_size // tp_world_size
        self.skip_bias_add = skip_bias_add
        self.weight = Parameter(torch.empty(output_size, self.input_size_per_partition, dtype=dtype, device=device))
        (init_method or torch.nn.init.xavier_normal_)(self.weight)
        if bias:
            self.bias = Parameter(torch.zeros(output_size, dtype=dtype, device=device))
        else:
            self.register_parameter("bias", None)

    def forward(self, input_):
        bias = self.bias if not self.skip_bias_add else None
        out = F.linear(input_, self.weight, bias)
        return out, (self.bias if self.skip_bias_add else None)
