"""flatten / unflatten (reference ``csrc/utils/py_ds_utils.cpp``: thin wrappers over torch utils)."""
from torch._utils import _flatten_dense_tensors as flatten  # noqa: F401
from torch._utils import _unflatten_dense_tensors as unflatten  # noqa: F401
