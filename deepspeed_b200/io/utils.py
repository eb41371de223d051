"""Serialisation helpers of the checkpoint writers (reference ``io/utils.py``)."""
from dataclasses import dataclass

import numpy as np
import torch

_DTYPES = {torch.float32: np.float32, torch.float64: np.float64, torch.float16: np.float16, torch.int8: np.int8,
           torch.uint8: np.uint8, torch.int16: np.int16, torch.int32: np.int32, torch.int64: np.int64,
           torch.bool: np.bool_}


@dataclass
class serialize_details:
    obj: object
    dtype: torch.dtype
    size: int
    nbytes: int


def required_minimum_torch_version(major_version=1, minor_version=12):
    parts = torch.__version__.split(".")
    return (int(parts[0]), int("".join(ch for ch in parts[1] if ch.isdigit()) or 0)) >= (major_version, minor_version)


def tensor_to_bytes(tensor):
    """The tensor's storage as a flat ``uint8`` tensor (a view: no copy)."""
    return tensor.detach().contiguous().view(-1).view(torch.uint8)


def bytes_to_tensor(buffer, dtype=torch.uint8):
    """A ``bytes`` / ``bytearray`` object as a tensor of ``dtype``."""
    flat = torch.frombuffer(bytearray(buffer), dtype=torch.uint8)
    return flat if dtype == torch.uint8 else flat.view(dtype)


def obj_serialization_details():
    """Accessor returning (object, dtype, element count, bytes) for a tensor or a typed / untyped storage - what a
    writer needs to stream a ``torch.save`` payload."""

    def details(obj):
        if torch.is_tensor(obj):
            return serialize_details(obj, obj.dtype, obj.numel(), obj.numel() * obj.element_size())
        untyped = obj.untyped() if hasattr(obj, "untyped") else obj
        dtype = getattr(obj, "dtype", torch.uint8)
        nbytes = untyped.nbytes()
        esz = torch.empty(0, dtype=dtype).element_size()
        return serialize_details(untyped, dtype, nbytes // esz, nbytes)

    return details
