"""Mixin for modules whose forward is replayed from a CUDA graph (reference ``model_implementations/features/cuda_graph.py``)."""


class CUDAGraph:

    def __init__(self, enable_cuda_graph=False):
        super().__init__()
        self.enable_cuda_graph = enable_cuda_graph

    def _create_cuda_graph(self, *inputs, **kwargs):
        raise NotImplementedError

    def _graph_replay(self, *inputs, **kwargs):
        raise NotImplementedError
