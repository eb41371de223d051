"""``MultiTensorApply`` (reference ``ops/adam/multi_tensor_apply.py``): calls a multi-tensor kernel entry point with the
chunk size fixed at construction. The fused optimizers of this framework (``csrc/multi_tensor_*.cu``) take tensor
lists directly; this wrapper is the call convention code written against the reference expects."""


class MultiTensorApply:

    def __init__(self, chunk_size):
        self.chunk_size = chunk_size

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
