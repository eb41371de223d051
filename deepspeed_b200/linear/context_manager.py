"""``deepspeed.linear.Init``: while active, ``nn.Linear`` constructs :class:`OptimizedLinear` so that a HuggingFace
``from_pretrained`` / ``from_config`` builds a LoRA / quantized model directly (reference ``linear/context_manager.py``)."""
import torch

from .optimized_linear import LoRAOptimizedLinear, OptimizedLinear

try:
    import transformers
except ImportError:  # pragma: no cover
    transformers = None


def init_lora(model):
    model.requires_grad_(False)
    for m in model.modules():
        if isinstance(m, LoRAOptimizedLinear):
            m.init_lora()


class Init:

    def __init__(self, lora_config=None, quant_config=None):
        self._orig_nn_linear = torch.nn.Linear
        self._orig_from_pretrained = transformers.AutoModelForCausalLM.from_pretrained if transformers else None
        self._orig_from_config = transformers.AutoModelForCausalLM.from_config if transformers else None
        self.lora_config, self.quant_config = lora_config, quant_config

    def __enter__(self):
        outer = self

        class OptLinearWrapper:
            _orig_nn_linear = outer._orig_nn_linear
            _lora_config, _quant_config = outer.lora_config, outer.quant_config

            def __new__(cls, *args, **kwargs):
                cls._lora_config.delay_lora_init = True
                kwargs["lora_config"], kwargs["quantization_config"] = cls._lora_config, cls._quant_config
                kwargs["linear_cls"] = cls._orig_nn_linear
                return OptimizedLinear(*args, **kwargs)

        def _model_init(model):
            if self.lora_config is not None:
                init_lora(model)
            return model

        def from_pretrained(*args, **kwargs):
            return _model_init(self._orig_from_pretrained(*args, **kwargs))

        def from_config(*args, **kwargs):
            return _model_init(self._orig_from_config(*args, **kwargs))

        torch.nn.Linear = OptLinearWrapper
        if transformers:
            transformers.AutoModelForCausalLM.from_pretrained = from_pretrained
            transformers.AutoModelForCausalLM.from_config = from_config
        return self

    def __exit__(self, *args, **kwargs):
        torch.nn.Linear = self._orig_nn_linear
        if transformers:
            transformers.AutoModelForCausalLM.from_pretrained = self._orig_from_pretrained
            transformers.AutoModelForCausalLM.from_config = self._orig_from_config
