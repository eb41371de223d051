"""``OptimizedLinear``: a linear layer with optional LoRA adapters, quantized frozen base weight and base-weight
sharding over the data-parallel group (reference ``linear/optimized_linear.py:18,76``)."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import comm as dist
from .config import LoRAConfig, QuantizationConfig
from .quantization import QuantizedLinear, QuantizedParameter


class OptimizedLinear(nn.Module):
    """Factory: plain ``nn.Linear`` without configs, :class:`QuantizedLinear` with only a quantization config,
    :class:`LoRAOptimizedLinear` when a LoRA config is given."""

    def __new__(cls, input_dim, output_dim, bias=False, lora_config: LoRAConfig = None,
                quantization_config: QuantizationConfig = None, device=None, dtype=torch.bfloat16,
                linear_cls=nn.Linear):
        if quantization_config is not None and not (lora_config is not None and lora_config.delay_lora_init):
            assert dtype == torch.bfloat16, "quantization requires bfloat16 weights"
        if lora_config is None and quantization_config is None:
            return nn.Linear(input_dim, output_dim, bias=bias, dtype=dtype, device=device)
        if lora_config is not None:
            return LoRAOptimizedLinear(input_dim, output_dim, bias=bias, lora_config=lora_config,
                                       quantization_config=quantization_config, device=device, dtype=dtype,
                                       linear_cls=linear_cls)
        return QuantizedLinear(input_dim, output_dim, bias=bias, quantization_config=quantization_config, dtype=dtype)


class LoRAOptimizedLinear(nn.Module):

    def __init__(self, input_dim, output_dim, bias=False, lora_config: LoRAConfig = None,
                 quantization_config: QuantizationConfig = None, device=None, dtype=torch.bfloat16,
                 linear_cls=nn.Linear):
        super().__init__()
        assert lora_config is not None, "DSOptimizedLinear requires a LoRA config"
        assert not bias, "bias=True is not supported by LoRAOptimizedLinear"
        self.input_dim, self.output_dim, self.bias = input_dim, output_dim, bias
        self.lora_config, self.quantization_config = lora_config, quantization_config
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu"
        self.device, self.linear_cls, self.dtype = device, linear_cls, dtype
        self.zero_shards = lora_config.base_weight_sharding
        self.sharded_weight_size = int(float(input_dim) // self.zero_shards)
        if self.zero_shards > 1:
            assert self.zero_shards == dist.get_world_size(), "base weight sharding is only supported across world size"
            w = torch.empty(output_dim * self.sharded_weight_size, dtype=dtype)
        else:
            w = torch.empty((output_dim, input_dim), dtype=dtype)
        nn.init.xavier_uniform_(w.view(self.sharded_weight_size, output_dim))
        w = nn.Parameter(w, requires_grad=False)
        if quantization_config is not None:
            assert dtype == torch.bfloat16, "only bfloat16 is supported when using quantization"
            self.weight = QuantizedParameter(w.data, quantization_config=quantization_config)
        else:
            self.weight = w
        self.disabled = False
        self._initialized = False
        if not lora_config.delay_lora_init:
            self.init_lora()

    def disable(self):
        """Module name is not a LoRA target: behave as a plain (full, unsharded) frozen linear."""
        self.disabled = True
        self.weight = nn.Parameter(torch.empty((self.output_dim, self.input_dim), dtype=self.dtype),
                                   requires_grad=False)

    def init_lora(self):
        if self.disabled:
            return
        if self.quantization_config is not None and not isinstance(self.weight, QuantizedParameter):
            self.weight = QuantizedParameter(self.weight.data, quantization_config=self.quantization_config)
        self._initialized = True
        self.weight.requires_grad = False
        self.weight.ds_optim_param = True  # ZeRO must not partition it again
        self.lora_scaling_factor = self.lora_config.lora_alpha / self.lora_config.lora_r
        r = self.lora_config.lora_r
        self.lora_weight_1 = self.linear_cls(self.input_dim, r, bias=self.bias, device=self.device, dtype=self.dtype)
        self.lora_weight_2 = self.linear_cls(r, self.output_dim, bias=self.bias, device=self.device, dtype=self.dtype)
        nn.init.kaiming_uniform_(self.lora_weight_1.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_weight_2.weight)
        self.lora_weight_1.weight.requires_grad = True
        self.lora_weight_2.weight.requires_grad = True

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        if not any(t in prefix for t in self.lora_config.target_mods):
            self.disable()
        elif self.zero_shards > 1:
            if not dist.is_initialized():
                raise RuntimeError("base weight sharding needs an initialised process group")
            n = self.output_dim * self.sharded_weight_size
            name = f"{prefix}weight"
            state_dict[name] = state_dict[name].flatten().narrow(0, dist.get_rank() * n, n)
        return super()._load_from_state_dict(state_dict, prefix, local_metadata, strict, missing_keys,
                                             unexpected_keys, error_msgs)

    def full_weight(self):
        """Dequantize (and un-offload) my shard, all-gather the shards into the [out, in] base weight."""
        w = self.weight
        offloaded = getattr(w, "ds_offload", False)
        if offloaded:
            w.offload(revert=True)
        local = w.dequantized() if isinstance(w, QuantizedParameter) else w.data
        if offloaded:
            w.offload()
        out = torch.empty(self.output_dim * self.input_dim, dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.reshape(-1).contiguous())
        return out.view(self.output_dim, self.input_dim)

    def forward(self, input_tensor):
        if self.disabled:
            return F.linear(input_tensor, self.weight)
        assert self._initialized, "init_lora was never called, please initialize before proceeding"
        if self.zero_shards > 1:
            with torch.no_grad():
                base = self.full_weight()
        elif self.quantization_config is not None:
            base = self.weight.dequantized()
        else:
            base = self.weight
        out = F.linear(input_tensor, base.to(input_tensor.dtype))
        lora = self.lora_weight_2(self.lora_weight_1(input_tensor))
        return out + self.lora_scaling_factor * lora
