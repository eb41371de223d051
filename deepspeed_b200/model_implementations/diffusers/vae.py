"""VAE of a diffusers pipeline: decode / encode / forward each behind their own CUDA graph (reference
``model_implementations/diffusers/vae.py``)."""
import torch

from ...compile.graph_capture import CUDAGraphRunner
from ..features.cuda_graph import CUDAGraph


class DSVAE(CUDAGraph, torch.nn.Module):

    def __init__(self, vae, enable_cuda_graph=True):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        self.vae = vae
        self.config = getattr(vae, "config", None)
        self.device, self.dtype = getattr(vae, "device", None), getattr(vae, "dtype", None)
        self.vae.requires_grad_(requires_grad=False)
        self._runners = {}

    def _run(self, name, fn, *inputs, **kwargs):
        if not (self.enable_cuda_graph and torch.cuda.is_available()):
            return fn(*inputs, **kwargs)
        if name not in self._runners:
            self._runners[name] = CUDAGraphRunner(lambda *a: fn(*a, **kwargs), *inputs)
        return self._runners[name](*inputs)

    def decode(self, *inputs, **kwargs):
        return self._run("decode", self.vae.decode, *inputs, **kwargs)

    def encode(self, *inputs, **kwargs):
        return self._run("encode", self.vae.encode, *inputs, **kwargs)

    def forward(self, *inputs, **kwargs):
        return self._run("forward", self.vae, *inputs, **kwargs)
