from ..adam.multi_tensor_apply import MultiTensorApply  # noqa: F401
