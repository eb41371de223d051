"""UNet of a diffusers pipeline behind CUDA-graph replay (reference ``model_implementations/diffusers/unet.py``)."""
import torch

from ...compile.graph_capture import CUDAGraphRunner
from ..features.cuda_graph import CUDAGraph


class DSUNet(CUDAGraph, torch.nn.Module):

    def __init__(self, unet, enable_cuda_graph=True):
        super().__init__(enable_cuda_graph=enable_cuda_graph)
        self.unet = unet
        self.in_channels = getattr(unet, "in_channels", None)
        self.device, self.dtype = getattr(unet, "device", None), getattr(unet, "dtype", None)
        self.config = getattr(unet, "config", None)
        self.fwd_count = 0
        self.unet.requires_grad_(requires_grad=False)
        self.unet.to(memory_format=torch.channels_last)
        self._runner = None

    def _forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, **kwargs):
        return self.unet(sample, timestamp, encoder_hidden_states, return_dict=return_dict, **kwargs)

    def forward(self, sample, timestamp, encoder_hidden_states, return_dict=True, **kwargs):
        if not (self.enable_cuda_graph and torch.cuda.is_available()):
            return self._forward(sample, timestamp, encoder_hidden_states, return_dict, **kwargs)
        if self._runner is None:
            self._runner = CUDAGraphRunner(lambda s, t, e: self._forward(s, t, e, return_dict, **kwargs), sample,
                                           timestamp, encoder_hidden_states)
        return self._runner(sample, timestamp, encoder_hidden_states)
