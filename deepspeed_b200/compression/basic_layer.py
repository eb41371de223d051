"""Compression-aware drop-in layers (reference ``compression/basic_layer.py``): weight / activation quantization (QAT
with straight-through estimators), sparse / row / head / channel pruning with L1 or learned top-k masks, and the
``fix_*`` helpers that bake masks into (optionally smaller) weights after training.

One mixin carries the technique state; Linear / Conv2d / Embedding / tensor-parallel linears derive from it."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn import init

from ..utils.logging import logger
from .utils import AsymQuantizer, BinaryQuantizer, SymQuantizer, TernaryQuantizer, TopKBinarizer


class QuantAct(nn.Module):
    """Static-range activation quantizer: tracks running min/max with momentum during training."""

    def __init__(self, act_range_momentum=0.95, quant_mode="symmetric"):
        super().__init__()
        self.act_range_momentum = act_range_momentum
        self.quant_mode = quant_mode
        self.act_function = SymQuantizer.apply if quant_mode == "symmetric" else AsymQuantizer.apply
        self.register_buffer("x_min_max", torch.zeros(2))

    def forward(self, x, num_bits, *args):
        if self.training:
            lo, hi = x.detach().min(), x.detach().max()
            if self.x_min_max[0] == self.x_min_max[1]:
                self.x_min_max[0], self.x_min_max[1] = lo, hi
            m = self.act_range_momentum
            self.x_min_max[0] = self.x_min_max[0] * m + lo * (1 - m)
            self.x_min_max[1] = self.x_min_max[1] * m + hi * (1 - m)
        return self.act_function(x, num_bits, self.x_min_max[0], self.x_min_max[1])


def _weight_quantizer(target_bits, quantization_type):
    if target_bits >= 3:
        return SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply
    assert quantization_type == "symmetric", "ternary/binary weight quantization is symmetric only"
    return TernaryQuantizer.apply if target_bits == 2 else BinaryQuantizer.apply


def _carry_quant_attrs(new, old):
    for a in ("start_bits", "target_bits", "q_period"):
        setattr(new, a, getattr(old, a, None))
    return new


class _CompressMixin:
    """Technique flags are flipped by the compression scheduler; masks are built by ``enable_*``."""

    def _init_compress_state(self):
        self.sparse_pruning_method = self.row_pruning_method = self.head_pruning_method = None
        self.channel_pruning_method = self.activation_quantization_method = None
        self.weight.start_bits = self.weight.target_bits = self.weight.q_period = None
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False
        self.sparse_pruning_enabled = self.row_pruning_enabled = self.head_pruning_enabled = False
        self.channel_pruning_enabled = self.activation_quantization_enabled = False

    # ---- quantization ---------------------------------------------------------------------
    def enable_weight_quantization(self, start_bits, target_bits, quantization_period,
                                   weight_quantization_enabled_in_forward, quantization_type, num_groups):
        self.weight.start_bits, self.weight.target_bits = start_bits, target_bits
        self.weight.q_period = quantization_period
        self.weight_quantization_enabled_in_forward = weight_quantization_enabled_in_forward
        if weight_quantization_enabled_in_forward:
            logger.warning("quantize_weight_in_forward: MoQ scheduling (bit annealing, eigenvalue) does not apply; "
                           "weights are quantized to target_bits in every forward")
            self.weight_quantizer = _weight_quantizer(target_bits, quantization_type)
            self.weight_quantize_num_groups = num_groups

    def fix_weight_quantization(self):
        self.weight.data = self.weight_quantizer(self.weight, self.weight.target_bits, None, None,
                                                 self.weight_quantize_num_groups).data
        self.weight_quantization_enabled_in_forward = False

    def enable_activation_quantization(self, bits, quantization_type, range_calibration):
        assert bits in (4, 8), "Only 4/8 bits activation quantization are supported for now"
        self.activation_quantization_bits = bits
        self.activation_quantization_method = f"{quantization_type}_{range_calibration}"
        if range_calibration == "static":
            self.activation_quantizer = QuantAct(quant_mode=quantization_type)
        else:
            self.activation_quantizer = SymQuantizer.apply if quantization_type == "symmetric" else AsymQuantizer.apply

    def _maybe_quantized_weight(self):
        if self.weight_quantization_enabled_in_forward and self.weight_quantization_enabled:
            return self.weight_quantizer(self.weight, self.weight.target_bits, None, None,
                                         self.weight_quantize_num_groups)
        return self.weight

    def _maybe_quantized_input(self, x):
        if not self.activation_quantization_enabled:
            return x
        groups = x.numel() // x.size(-1) if "dynamic" in self.activation_quantization_method else 1
        return self.activation_quantizer(x, self.activation_quantization_bits, None, None, groups)

    # ---- sparse pruning (both Linear and Conv2d) ----------------------------------------------
    def enable_sparse_pruning(self, ratio, method):
        self.sparse_pruning_ratio, self.sparse_pruning_method = ratio, method
        mask = None
        if method == "l1":
            mask = TopKBinarizer.apply(self.weight.data.abs(), ratio, False).view(self.weight.size())
        elif method == "topk":
            self.sparse_mask_scores = nn.Parameter(torch.empty_like(self.weight.data))
            init.kaiming_uniform_(self.sparse_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError(method)
        self.register_buffer("sparse_pruning_mask", mask)

    def fix_sparse_pruning_helper(self):
        self.weight.data = self.weight.data * self.get_mask("sparse")
        del self.sparse_pruning_mask
        if self.sparse_pruning_method == "topk":
            del self.sparse_mask_scores
        self.sparse_pruning_method, self.sparse_pruning_enabled = None, False

    def _mask(self, kind, method, buffer_name, scores_name, ratio):
        if method == "l1":
            return getattr(self, buffer_name).to(self.weight.device)
        if method == "topk":
            return TopKBinarizer.apply(getattr(self, scores_name), ratio, False)
        raise NotImplementedError(f"{kind} pruning method {method}")


class LinearLayer_Compress(nn.Linear, _CompressMixin):

    def __init__(self, *kargs, bias=True):
        super().__init__(*kargs, bias=bias)
        self._init_compress_state()

    def extra_repr(self):
        return (f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, "
                f"sparse pruning={self.sparse_pruning_method is not None}, "
                f"row pruning={self.row_pruning_method is not None}, "
                f"head pruning={self.head_pruning_method is not None}, "
                f"activation quantization={self.activation_quantization_method is not None}, "
                f"weight_quantization={self.weight.target_bits}")

    def enable_row_pruning(self, ratio, method):
        self.row_pruning_ratio, self.row_pruning_method = ratio, method
        mask = None
        if method == "l1":
            mask = TopKBinarizer.apply(torch.linalg.norm(self.weight.data, ord=1, dim=1), ratio, False).view(-1, 1)
        elif method == "topk":
            self.row_mask_scores = nn.Parameter(torch.empty(self.weight.size(0), 1, device=self.weight.device))
            init.kaiming_uniform_(self.row_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError(method)
        self.register_buffer("row_pruning_mask", mask)

    def enable_head_pruning(self, ratio, method, num_heads):
        if method != "topk":
            raise NotImplementedError("head pruning supports method 'topk' only")
        self.num_heads, self.head_pruning_ratio, self.head_pruning_method = num_heads, ratio, method
        # applied to the attention output projection: one score per head (= a block of input columns)
        self.head_pruning_scores = nn.Parameter(torch.empty(1, num_heads, device=self.weight.device))
        init.kaiming_uniform_(self.head_pruning_scores, a=math.sqrt(5))

    def get_mask(self, pruning_type="row"):
        if pruning_type == "sparse":
            return self._mask("sparse", self.sparse_pruning_method, "sparse_pruning_mask", "sparse_mask_scores",
                              getattr(self, "sparse_pruning_ratio", None))
        if pruning_type == "row":
            return self._mask("row", self.row_pruning_method, "row_pruning_mask", "row_mask_scores",
                              getattr(self, "row_pruning_ratio", None))
        if pruning_type == "head":
            if self.head_pruning_method == "topk":
                return TopKBinarizer.apply(self.head_pruning_scores, self.head_pruning_ratio, False)
        raise NotImplementedError(pruning_type)

    def _replace_weight(self, data):
        self.weight = _carry_quant_attrs(nn.Parameter(data), self.weight)

    def fix_row_col_pruning_helper(self, mask=None, dim_reduction=False):
        """Own mask: zero (or physically drop) output rows. Foreign ``mask`` (from the preceding layer): drop the
        matching input columns."""
        if mask is None:
            mask = self.get_mask("row").bool()
            if dim_reduction:
                self._replace_weight(self.weight.data[mask.view(-1), :])
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[mask.view(-1)])
                self.out_features = self.weight.size(0)
            else:
                self.weight.data = self.weight.data * mask.view(-1, 1)
                if self.bias is not None:
                    self.bias.data = self.bias.data * mask.view(-1)
            del self.row_pruning_mask
            if self.row_pruning_method == "topk":
                del self.row_mask_scores
            self.row_pruning_method = None
        else:
            self._replace_weight(self.weight.data[:, mask.view(-1)])
            self.in_features = self.weight.size(1)
            mask = None
        self.row_pruning_enabled = False
        return mask

    def fix_head_pruning_helper(self, mask=None, num_heads=None, dim_reduction=False):
        """Own mask (output projection): zero/drop head-sized column blocks. Foreign mask (Q/K/V): drop the rows."""
        num_heads = num_heads or self.num_heads
        if mask is None:
            if self.head_pruning_method != "topk":
                raise NotImplementedError
            mask = self.get_mask("head").bool()
            if dim_reduction:
                rows = self.weight.size(0)
                self._replace_weight(self.weight.data.t().reshape(num_heads, -1)[mask.view(-1), :].reshape(-1, rows).t())
                self.in_features = self.weight.size(1)
            else:
                shape = self.weight.size()
                self.weight.data = (self.weight.data.t().reshape(self.num_heads, -1) * mask.view(-1, 1)).reshape(
                    shape[1], shape[0]).t()
            del self.head_pruning_scores
            self.head_pruning_method = None
        else:
            cols = self.weight.size(1)
            self._replace_weight(self.weight.data.reshape(num_heads, -1)[mask.view(-1), :].reshape(-1, cols))
            if self.bias is not None:
                self.bias = nn.Parameter(self.bias.data.reshape(num_heads, -1)[mask.view(-1), :].reshape(-1))
            self.out_features = self.weight.size(0)
        self.head_pruning_enabled = False
        return mask

    def head_pruning_reshape(self, w, mask):
        shape = w.shape
        return (w.t().reshape(self.num_heads, -1) * mask.view(-1, 1)).reshape(shape[1], shape[0]).t()

    def forward(self, input, skip_bias_add=False):
        weight, bias = self._maybe_quantized_weight(), self.bias
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            weight = weight * self.get_mask("sparse").view(self.weight.size())
        if self.row_pruning_enabled and self.row_pruning_method:
            mask = self.get_mask("row")
            weight = weight * mask.view(-1, 1)
            if bias is not None:
                bias = bias * mask.view(-1)
        if self.head_pruning_enabled and self.head_pruning_method:
            weight = self.head_pruning_reshape(weight, self.get_mask("head"))
        input = self._maybe_quantized_input(input)
        if skip_bias_add:
            return F.linear(input, weight, None), bias
        return F.linear(input, weight, bias)


class Embedding_Compress(nn.Embedding, _CompressMixin):

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self.weight.start_bits = self.weight.target_bits = self.weight.q_period = None
        self.weight_quantization_enabled_in_forward = False
        self.weight_quantization_enabled = False

    def extra_repr(self):
        return (f"num_embeddings={self.num_embeddings}, embedding_dim={self.embedding_dim}, "
                f"weight_quantization={self.weight.target_bits}")

    def enable_weight_quantization(self, start_bits, target_bits, quantization_period,
                                   weight_quantization_enabled_in_forward, quantization_type, num_groups):
        super().enable_weight_quantization(start_bits, target_bits, quantization_period,
                                           weight_quantization_enabled_in_forward, quantization_type, num_groups)
        if weight_quantization_enabled_in_forward:
            # embeddings are quantized token-wise: one group per row
            self.weight_quantize_num_groups = self.weight.size(0)

    def forward(self, input):
        return F.embedding(input, self._maybe_quantized_weight(), self.padding_idx, self.max_norm, self.norm_type,
                           self.scale_grad_by_freq, self.sparse)


class Conv2dLayer_Compress(nn.Conv2d, _CompressMixin):

    def __init__(self, *kargs):
        super().__init__(*kargs)
        self._init_compress_state()

    def enable_channel_pruning(self, ratio, method):
        self.channel_pruning_ratio, self.channel_pruning_method = ratio, method
        mask = None
        if method == "l1":
            norm = self.weight.data.abs().sum(dim=(1, 2, 3))
            mask = TopKBinarizer.apply(norm, ratio, False).view(-1, 1, 1, 1)
        elif method == "topk":
            self.channel_mask_scores = nn.Parameter(torch.empty(self.weight.size(0), 1, 1, 1,
                                                                device=self.weight.device))
            init.kaiming_uniform_(self.channel_mask_scores, a=math.sqrt(5))
        else:
            raise NotImplementedError(method)
        self.register_buffer("channel_pruning_mask", mask)

    def get_mask(self, pruning_type="sparse"):
        if pruning_type == "sparse":
            return self._mask("sparse", self.sparse_pruning_method, "sparse_pruning_mask", "sparse_mask_scores",
                              getattr(self, "sparse_pruning_ratio", None))
        if pruning_type == "channel":
            return self._mask("channel", self.channel_pruning_method, "channel_pruning_mask", "channel_mask_scores",
                              getattr(self, "channel_pruning_ratio", None))
        raise NotImplementedError(pruning_type)

    def fix_channel_pruning_helper(self, mask=None, dim_reduction=False):
        if mask is None:
            if self.channel_pruning_method not in ("l1", "topk"):
                raise NotImplementedError
            mask = self.get_mask("channel").bool()
            if dim_reduction:
                self.weight = _carry_quant_attrs(nn.Parameter(self.weight.data[mask.view(-1), ...]), self.weight)
                if self.bias is not None:
                    self.bias = nn.Parameter(self.bias.data[mask.view(-1)])
                self.out_channels = self.weight.size(0)
            else:
                self.weight.data = self.weight.data * mask.view(-1, 1, 1, 1)
                if self.bias is not None:
                    self.bias.data = self.bias.data * mask.view(-1)
            del self.channel_pruning_mask
            if self.channel_pruning_method == "topk":
                del self.channel_mask_scores
            self.channel_pruning_method = None
        else:
            self.weight = _carry_quant_attrs(nn.Parameter(self.weight.data[:, mask.view(-1), ...]), self.weight)
            self.in_channels = self.weight.size(1)
            mask = None
        self.channel_pruning_enabled = False
        return mask

    def forward(self, input):
        weight, bias = self._maybe_quantized_weight(), self.bias
        if self.sparse_pruning_enabled and self.sparse_pruning_method:
            weight = weight * self.get_mask("sparse").view(self.weight.size())
        if self.channel_pruning_enabled:
            mask = self.get_mask("channel")
            weight = weight * mask.view(-1, 1, 1, 1)
            if bias is not None:
                bias = bias * mask.view(-1)
        input = self._maybe_quantized_input(input)
        return F.conv2d(input, weight, bias, self.stride, self.padding, self.dilation, self.groups)


class BNLayer_Compress(nn.BatchNorm2d):
    """BatchNorm following a channel-pruned conv: shrink (or zero) its statistics with the conv's mask."""

    def fix_channel_pruning_helper(self, mask, dim_reduction=True):
        keep = mask.view(-1).bool()
        if dim_reduction:
            self.weight = nn.Parameter(self.weight.data[keep])
            self.bias = nn.Parameter(self.bias.data[keep])
            self.running_mean = self.running_mean[keep]
            self.running_var = self.running_var[keep]
            self.num_features = self.weight.size(0)
        else:
            self.weight.data *= keep
            self.bias.data *= keep
            self.running_mean *= keep
            self.running_var *= keep
        return None


# ---- tensor-parallel (Megatron mpu) variants -------------------------------------------------------------------
def _mp_size(mpu):
    return mpu.get_model_parallel_world_size()


class _CopyToModelParallelRegion(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return input_

    @staticmethod
    def backward(ctx, grad_output):
        if _mp_size(ctx.mpu) > 1:
            grad_output = grad_output.contiguous()
            torch.distributed.all_reduce(grad_output, group=ctx.mpu.get_model_parallel_group())
        return grad_output, None


class _ReduceFromModelParallelRegion(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, mpu):
        if _mp_size(mpu) > 1:
            input_ = input_.contiguous()
            torch.distributed.all_reduce(input_, group=mpu.get_model_parallel_group())
        return input_

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output, None


def _split_last(t, mpu):
    ws = _mp_size(mpu)
    return t if ws == 1 else t.chunk(ws, dim=-1)[mpu.get_model_parallel_rank()].contiguous()


def _gather_last(t, mpu):
    ws = _mp_size(mpu)
    if ws == 1:
        return t
    parts = [torch.empty_like(t) for _ in range(ws)]
    torch.distributed.all_gather(parts, t.contiguous(), group=mpu.get_model_parallel_group())
    parts[mpu.get_model_parallel_rank()] = t
    return torch.cat(parts, dim=-1)


class _ScatterToModelParallelRegion(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return _split_last(input_, mpu)

    @staticmethod
    def backward(ctx, grad_output):
        return _gather_last(grad_output, ctx.mpu), None


class _GatherFromModelParallelRegion(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input_, mpu):
        ctx.mpu = mpu
        return _gather_last(input_, mpu)

    @staticmethod
    def backward(ctx, grad_output):
        return _split_last(grad_output, ctx.mpu), None


def copy_to_model_parallel_region(input_, mpu):
    return _CopyToModelParallelRegion.apply(input_, mpu)


def reduce_from_model_parallel_region(input_, mpu):
    return _ReduceFromModelParallelRegion.apply(input_, mpu)


def scatter_to_model_parallel_region(input_, mpu):
    return _ScatterToModelParallelRegion.apply(input_, mpu)


def gather_from_model_parallel_region(input_, mpu):
    return _GatherFromModelParallelRegion.apply(input_, mpu)


class ColumnParallelLinear_Compress(LinearLayer_Compress):

    def __init__(self, mpu, input_size, output_size, bias=True, gather_output=True, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.gather_output, self.skip_bias_add = gather_output, skip_bias_add
        ws = _mp_size(mpu)
        assert output_size % ws == 0
        self.output_size_per_partition = output_size // ws
        super().__init__(input_size, self.output_size_per_partition, bias=bias)

    def forward(self, input_):
        out, bias = super().forward(copy_to_model_parallel_region(input_, self.mpu), True)
        if self.gather_output:
            out = gather_from_model_parallel_region(out, self.mpu)
        if self.skip_bias_add:
            return out, bias
        return out if bias is None else out + (gather_from_model_parallel_region(bias, self.mpu)
                                               if self.gather_output else bias)


class RowParallelLinear_Compress(LinearLayer_Compress):

    def __init__(self, mpu, input_size, output_size, bias=True, input_is_parallel=False, skip_bias_add=False):
        self.mpu = mpu
        self.input_size, self.output_size = input_size, output_size
        self.input_is_parallel, self.skip_bias_add = input_is_parallel, skip_bias_add
        ws = _mp_size(mpu)
        assert input_size % ws == 0
        self.input_size_per_partition = input_size // ws
        super().__init__(self.input_size_per_partition, output_size, bias=bias)

    def forward(self, input_):
        x = input_ if self.input_is_parallel else scatter_to_model_parallel_region(input_, self.mpu)
        partial, bias = super().forward(x, True)
        out = reduce_from_model_parallel_region(partial, self.mpu)
        if self.skip_bias_add:
            return out, bias
        return out if bias is None else out + bias
