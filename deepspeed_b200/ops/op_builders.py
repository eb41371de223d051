"""Builder shims with the reference's names (``FusedAdamBuilder().load()`` ...). Everything is prebuilt into
``libdsb200.so`` / ``libdsb200_cpu.so`` by :mod:`deepspeed_b200.ops.builder`; ``load()`` returns the Python module
that exposes the op."""
import importlib


class _Builder:
    NAME = "op"
    MODULE = None

    def is_compatible(self, verbose=False):
        return True

    def load(self, verbose=False):
        return importlib.import_module(self.MODULE)

    def builder(self):
        return self

    def absolute_name(self):
        return f"deepspeed_b200.ops.{self.NAME}"


def _mk(name, module):
    return type(name, (_Builder, ), {"NAME": name.replace("Builder", "").lower(), "MODULE": module})


FusedAdamBuilder = _mk("FusedAdamBuilder", "deepspeed_b200.ops.optimizers")
FusedLambBuilder = _mk("FusedLambBuilder", "deepspeed_b200.ops.optimizers")
FusedLionBuilder = _mk("FusedLionBuilder", "deepspeed_b200.ops.optimizers")
CPUAdamBuilder = _mk("CPUAdamBuilder", "deepspeed_b200.ops.cpu_optimizers")
CPUAdagradBuilder = _mk("CPUAdagradBuilder", "deepspeed_b200.ops.cpu_optimizers")
CPULionBuilder = _mk("CPULionBuilder", "deepspeed_b200.ops.cpu_optimizers")
AsyncIOBuilder = _mk("AsyncIOBuilder", "deepspeed_b200.ops.aio")
QuantizerBuilder = _mk("QuantizerBuilder", "deepspeed_b200.ops.quantizer")
FPQuantizerBuilder = _mk("FPQuantizerBuilder", "deepspeed_b200.ops.fp_quantizer")
UtilsBuilder = _mk("UtilsBuilder", "deepspeed_b200.ops.utils_op")
TransformerBuilder = _mk("TransformerBuilder", "deepspeed_b200.ops.transformer")
RandomLTDBuilder = _mk("RandomLTDBuilder", "deepspeed_b200.ops.random_ltd")
SparseAttnBuilder = _mk("SparseAttnBuilder", "deepspeed_b200.ops.sparse_attention")
EvoformerAttnBuilder = _mk("EvoformerAttnBuilder", "deepspeed_b200.ops.deepspeed4science")
InferenceBuilder = _mk("InferenceBuilder", "deepspeed_b200.inference.engine")
SpatialInferenceBuilder = _mk("SpatialInferenceBuilder", "deepspeed_b200.ops.spatial")
PinMemoryBuilder = _mk("PinMemoryBuilder", "deepspeed_b200.ops.pin_memory")
GDSBuilder = _mk("GDSBuilder", "deepspeed_b200.ops.gds")
StochasticTransformerBuilder = _mk("StochasticTransformerBuilder", "deepspeed_b200.ops.transformer")
RaggedOpsBuilder = _mk("RaggedOpsBuilder", "deepspeed_b200.inference.v2.ragged")
DeepCompileBuilder = _mk("DeepCompileBuilder", "deepspeed_b200.compile")
InferenceCoreBuilder = _mk("InferenceCoreBuilder", "deepspeed_b200.ops.fused")  # norms / activations / rotary kernels
InferenceCutlassBuilder = _mk("InferenceCutlassBuilder", "deepspeed_b200.ops.kernels")  # wo_gemm, grouped GEMM
RaggedUtilsBuilder = _mk("RaggedUtilsBuilder", "deepspeed_b200.inference.v2.ragged")  # csrc/cpu/ragged_host.cpp
ALL_OPS = {k: v for k, v in globals().items() if k.endswith("Builder") and k != "_Builder"}

# base-class names of the reference's op_builder package (``isinstance(b, OpBuilder)`` checks in user code)
OpBuilder = _Builder
CUDAOpBuilder = _Builder
TorchCPUOpBuilder = _Builder


def get_default_compute_capabilities():
    return "10.0a"
