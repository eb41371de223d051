"""Model families implemented on the fused ops: GPT-2, Llama (2/3), Mixtral (MoE)."""
from .gpt2 import GPT2Config, GPT2LMHeadModel
from .llama import LlamaConfig, LlamaForCausalLM, LlamaModel


def __getattr__(name):
    if name in ("MixtralConfig", "MixtralForCausalLM"):
        from . import mixtral
        return getattr(mixtral, name)
    raise AttributeError(name)
