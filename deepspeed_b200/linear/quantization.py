"""``QuantizedParameter`` / ``QuantizedLinear`` (reference ``linear/quantization.py:18,129``): a frozen weight stored
as block-scaled FP8/FP6/FP4 and dequantized on use. The payload is the parameter's data; scales ride along as an
attribute. Quantization happens when the parameter first lands on a CUDA device (or immediately on CPU with the
pure-PyTorch path), through ``ops.fp_quantizer.FP_Quantize`` (sm_100a ``minifloat_quant`` kernel on GPU)."""
import copy

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops.fp_quantizer import FP_Quantize
from .config import QuantizationConfig


class QuantizedParameter(nn.Parameter):

    def __new__(cls, data=None, requires_grad=False, quantization_config=None, quantizer=None):
        if requires_grad:
            raise ValueError("requires_grad=True is not supported with QuantizedParameter")
        if data is None:
            data = torch.empty(0)
        self = torch.Tensor._make_subclass(cls, data, requires_grad)
        self.quantization_config = QuantizationConfig() if quantization_config is None else quantization_config
        self.quantizer = quantizer if quantizer is not None else FP_Quantize(quantization_config=self.quantization_config)
        self._q = None
        self._ensure_quantized(self)
        return self

    def _ensure_quantized(self, tensor):
        # quantize once, when the floating-point data is resident where it will be used
        if self._q is None and tensor.is_floating_point() and tensor.numel() > 0 and \
                (tensor.is_cuda or not torch.cuda.is_available()):
            g = self.quantization_config.group_size
            assert tensor.numel() % g == 0, f"numel {tensor.numel()} not divisible by group_size {g}"
            with torch.no_grad():
                self._q = self.quantizer.quantize(tensor.data, q_bits=self.quantization_config.q_bits,
                                                  q_mantisa_bits=self.quantization_config.mantissa_bits)
                self._orig_shape, self._orig_dtype = tensor.shape, tensor.dtype
                self.data = self._q.payload.view(torch.uint8) if self._q.payload.dtype != torch.bfloat16 \
                    else self._q.payload

    def dequantized(self):
        """Floating-point copy of the weight (bf16 by default), on the payload's device."""
        if self._q is None:
            return self.data
        q = self._q
        if q.payload.device != self.data.device:  # moved after quantization
            q.payload = self.data.view(q.payload.dtype)
            q.scale = q.scale.to(self.data.device)
        with torch.no_grad():
            return self.quantizer.dequantize(q)

    def offload(self, revert=False):
        dev = (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else
               torch.device("cpu")) if revert else torch.device("cpu")
        self.data = self.data.to(dev)
        if self._q is not None:
            self._q.payload = self.data.view(self._q.payload.dtype)
            self._q.scale = self._q.scale.to(dev)

    def __getstate__(self):
        state = self.__dict__.copy()
        state["data"] = self.data
        state["requires_grad"] = self.requires_grad
        return state

    def __setstate__(self, state):
        self.quantizer = state["quantizer"]
        self.quantization_config = state["quantization_config"]
        self._q = state.get("_q")
        self.data = state["data"]
        self.requires_grad = state["requires_grad"]

    def __deepcopy__(self, memo):
        state = self.__getstate__()
        new = type(self).__new__(type(self), state["data"], quantization_config=state["quantization_config"],
                                 quantizer=copy.deepcopy(state["quantizer"]))
        new._q = copy.deepcopy(self._q)
        if self._q is not None:
            new._orig_shape, new._orig_dtype = self._orig_shape, self._orig_dtype
        return new

    def __copy__(self):
        state = self.__getstate__()
        new = type(self).__new__(type(self), state["data"], quantization_config=state["quantization_config"],
                                 quantizer=state["quantizer"])
        new._q = self._q
        return new

    def cuda(self, device=None, non_blocking=False):
        return self.to(device="cuda" if device is None else device, non_blocking=non_blocking)

    def to(self, *args, **kwargs):
        moved = super().to(*args, **kwargs)
        out = torch.Tensor._make_subclass(QuantizedParameter, moved.data if isinstance(moved, torch.Tensor) else moved,
                                          False)
        out.quantization_config, out.quantizer, out._q = self.quantization_config, self.quantizer, self._q
        if self._q is not None:
            out._orig_shape, out._orig_dtype = self._orig_shape, self._orig_dtype
            out._q = copy.copy(self._q)
            out._q.payload = out.data.view(self._q.payload.dtype)
            out._q.scale = self._q.scale.to(out.data.device)
        else:
            out._ensure_quantized(out)
        return out


class QuantizedLinear(nn.Linear):
    """Linear whose weight is a :class:`QuantizedParameter`; bias is not supported (as in the reference)."""

    def __init__(self, input_dim, output_dim, bias=False, quantization_config=None, dtype=torch.bfloat16):
        super().__init__(input_dim, output_dim, bias=bias, dtype=dtype)
        assert dtype == torch.bfloat16, "currently only supports bfloat16 dtype"
        self.weight = QuantizedParameter(self.weight.data, quantization_config=quantization_config)

    def _apply(self, fn, recurse=True):
        # nn.Module._apply assigns ``param.data = fn(param)``, which would drop the quantization metadata (and, for a
        # weight still waiting to be quantized on its first CUDA residency, skip the quantization): convert the weight
        # through QuantizedParameter.to instead
        w = self._parameters.pop("weight")
        try:
            super()._apply(fn)
            with torch.no_grad():
                nw = fn(w)
        finally:
            self._parameters["weight"] = w
        if not isinstance(nw, QuantizedParameter):
            nw = QuantizedParameter(nw.data, quantization_config=w.quantization_config, quantizer=w.quantizer)
        self._parameters["weight"] = nw
        return self

    def forward(self, input):
        w = self.weight
        q = getattr(w, "_q", None)
        if q is not None and input.is_cuda and not torch.is_grad_enabled():
            # decode-sized inputs: the quantized weights cross HBM once and are expanded in registers
            # (csrc/wo_gemm.cu); nothing is dequantized to memory
            from ..ops import kernels as K
            x2 = input.reshape(-1, input.shape[-1])
            N, Kd = w._orig_shape[0], w._orig_shape[1]
            if K.wo_gemm_supported(x2, w.data, q.group_size, q.q_bits, Kd) and q.scale.is_cuda:
                y = K.wo_gemm(x2, w.data, q.scale, N, Kd, q.group_size, q.q_bits, self.bias)
                return y.view(*input.shape[:-1], N)
        return F.linear(input, self.weight.dequantized(), self.bias)
