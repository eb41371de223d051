"""Memory-efficient linear for ZeRO-3 (reference ``runtime/zero/linear.py:80-148``): saves only the *handles* of input
and weight, so a partitioned weight is not kept gathered between forward and backward; backward re-reads ``weight``
(which the ZeRO-3 pre-backward hook has re-gathered)."""
import math

import torch
from torch import Tensor
from torch.nn import init
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter


class LinearFunctionForZeroStage3(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, weight, bias=None):
        ctx.save_for_backward(input, weight, bias)
        if input.dim() == 2 and bias is not None:
            return torch.addmm(bias, input, weight.t())
        out = input.matmul(weight.t())
        return out + bias if bias is not None else out

    @staticmethod
    def backward(ctx, grad_output):
        input, weight, bias = ctx.saved_tensors
        grad_input = grad_weight = grad_bias = None
        if ctx.needs_input_grad[0]:
            grad_input = grad_output.matmul(weight)
        if ctx.needs_input_grad[1]:
            go = grad_output.reshape(-1, grad_output.shape[-1])
            grad_weight = go.t().matmul(input.reshape(-1, input.shape[-1]))
        if bias is not None and ctx.needs_input_grad[2]:
            grad_bias = grad_output.reshape(-1, grad_output.shape[-1]).sum(0)
        return grad_input, grad_weight, grad_bias


def zero3_linear_wrap(input, weight, bias=None):
    return LinearFunctionForZeroStage3.apply(input, weight, bias) if bias is not None else \
        LinearFunctionForZeroStage3.apply(input, weight)


class LinearModuleForZeroStage3(Module):
    __constants__ = ["in_features", "out_features"]

    def __init__(self, in_features: int, out_features: int, bias: bool = True) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = Parameter(torch.empty(out_features, in_features))
        self.bias = Parameter(torch.empty(out_features)) if bias else None
        if not bias:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self) -> None:
        init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in, _ = init._calculate_fan_in_and_fan_out(self.weight)
            bound = 1 / math.sqrt(fan_in)
            init.uniform_(self.bias, -bound, bound)

    def forward(self, input: Tensor) -> Tensor:
        return LinearFunctionForZeroStage3.apply(input, self.weight, self.bias)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"
