"""CLIP text encoder of a diffusers pipeline behind CUDA-graph replay (reference
``model_implementations/transformers/clip_encoder.py``). Two graphs alternate because the pipeline calls the encoder
twice per image (prompt and negative prompt) with different static inputs."""
import torch

from ...compile.graph_capture import CUDAGraphRunner
from ..features.cuda_graph import CUDAGraph


class DSClipEncoder(CUDAGraph, torch.nn.Module):

    def __init__(self, enc, enable_cuda_graph=False):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        if hasattr(enc, "text_model") and hasattr(enc.text_model, "_build_causal_attention_mask"):
            enc.text_model._build_causal_attention_mask = self._build_causal_attention_mask
        self.enc = enc
        self.device, self.dtype = getattr(enc, "device", None), getattr(enc, "dtype", None)
        self.config = getattr(enc, "config", None)
        self._runners = [None, None]
        self.iter = 0

    def _build_causal_attention_mask(self, bsz, seq_len, dtype):
        mask = torch.full((bsz, seq_len, seq_len), torch.finfo(dtype).min, dtype=dtype,
                          device="cuda" if torch.cuda.is_available() else "cpu")
        return mask.triu_(1).unsqueeze(1)

    def _forward(self, *inputs, **kwargs):
        return self.enc(*inputs, **kwargs)

    def forward(self, *inputs, **kwargs):
        if not (self.enable_cuda_graph and torch.cuda.is_available()):
            return self.enc(*inputs, **kwargs)
        i = self.iter
        self.iter = (self.iter + 1) % 2
        if self._runners[i] is None:
            keys = list(kwargs)
            self._runners[i] = (CUDAGraphRunner(lambda *a: self._forward(*a[:len(inputs)],
                                                                         **dict(zip(keys, a[len(inputs):]))),
                                                *inputs, *[kwargs[k] for k in keys]), keys)
        runner, keys = self._runners[i]
        return runner(*inputs, *[kwargs[k] for k in keys])
