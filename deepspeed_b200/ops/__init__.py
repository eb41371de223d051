"""Operators (reference ``deepspeed/ops/__init__.py``): optimizer ops, transformer layers, sparse attention, quantizers;
sub-packages are imported on first use (the CUDA library is loaded lazily)."""
import importlib

_SUBPACKAGES = ("adam", "adagrad", "lamb", "lion", "sparse_attention", "transformer", "fp_quantizer", "aio", "gds",
                "quantizer", "random_ltd", "spatial", "deepspeed4science", "op_builder", "pin_memory", "kernels",
                "attention", "fused", "grouped_gemm", "optimizers", "cpu_optimizers", "muon", "builder", "op_builders",
                "cpu_lib", "utils_op")


def __getattr__(name):
    if name in _SUBPACKAGES:
        return importlib.import_module(f"{__name__}.{name}")
    if name in ("DeepSpeedTransformerLayer", "DeepSpeedTransformerConfig"):
        return getattr(importlib.import_module(f"{__name__}.transformer"), name)
    if name == "__compatible_ops__":
        from .op_builders import ALL_OPS
        return {cls.NAME if hasattr(cls, "NAME") else name: cls().is_compatible() for name, cls in ALL_OPS.items()}
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
