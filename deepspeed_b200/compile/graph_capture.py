"""CUDA-graph capture of launch-bound step functions ("CUDA streams and graphs instead of a tracing compiler";
reference analogue: ``runtime/hybrid_engine_graph.py`` / inference ``enable_cuda_graph``).

``CUDAGraphRunner(fn, *example_inputs)`` warms ``fn`` up on a side stream, captures it once with static input/output
buffers, and afterwards ``runner(*inputs)`` copies the inputs into the static buffers and replays the graph."""
import torch


class CUDAGraphRunner:

    def __init__(self, fn, *example_inputs, warmup=3, pool=None):
        assert torch.cuda.is_available(), "CUDA graphs need a CUDA device"
        self.fn = fn
        self.static_inputs = [x.clone() if torch.is_tensor(x) else x for x in example_inputs]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn(*self.static_inputs)
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, pool=pool):
            self.static_outputs = fn(*self.static_inputs)

    def __call__(self, *inputs):
        for dst, src in zip(self.static_inputs, inputs):
            if torch.is_tensor(dst):
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_outputs
