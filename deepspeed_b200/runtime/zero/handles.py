"""Work handles and small records of the ZeRO-3 parameter machinery under the reference's names
(``runtime/zero/partition_parameters.py:667-850``, ``partitioned_param_coordinator.py:33-76``).

The unit coordinator of this framework gathers whole flat units and waits on the collective's own work object; these
classes give code written against the reference the same ``handle.wait()`` objects for single parameters, coalesced
groups, already-available parameters and reductions."""
from collections import OrderedDict
from dataclasses import dataclass
from enum import Enum
from typing import Any, List

import torch

from .param_status import ZeroParamStatus


def _mark(params, status):
    for p in params:
        if hasattr(p, "ds_status"):
            p.ds_status = status


class NoGatherHandle:
    """The parameter was already whole (world size one or persistent): nothing to wait for."""

    def __init__(self, param):
        if getattr(param, "ds_status", ZeroParamStatus.INFLIGHT) not in (ZeroParamStatus.INFLIGHT, "INFLIGHT"):
            raise RuntimeError(f"expected param {getattr(param, 'ds_id', '?')} to be inflight")
        self.__param = param

    def wait(self, **kwargs):
        _mark([self.__param], ZeroParamStatus.AVAILABLE)


class NoGatherCoalescedHandle:

    def __init__(self, params):
        self.__params = list(params)
        self.__complete = False
        for p in self.__params:
            if getattr(p, "ds_status", ZeroParamStatus.INFLIGHT) not in (ZeroParamStatus.INFLIGHT, "INFLIGHT"):
                raise RuntimeError(f"expected param {getattr(p, 'ds_id', '?')} to not be available")

    def wait(self, **kwargs):
        if not self.__complete:
            _mark(self.__params, ZeroParamStatus.AVAILABLE)
            self.__complete = True


class AllGatherHandle:

    def __init__(self, handle, param, quantization=None):
        if getattr(param, "ds_status", ZeroParamStatus.INFLIGHT) not in (ZeroParamStatus.INFLIGHT, "INFLIGHT"):
            raise RuntimeError(f"expected param {getattr(param, 'ds_id', '?')} to be inflight")
        self.__handle, self.__param, self.__quantization = handle, param, quantization

    def wait(self, handle_dependency=True):
        if self.__handle is not None and hasattr(self.__handle, "wait"):
            self.__handle.wait()
        q = self.__quantization
        if q is not None and q.backend is not None:
            self.__param.data = q.backend.dequantize(q.quantized_param, q.scale_buffer).to(self.__param.device)
        _mark([self.__param], ZeroParamStatus.AVAILABLE)


class AllGatherCoalescedHandle:
    """One collective brought a flat buffer with every rank's partitions of several parameters: ``wait()`` completes it
    and points each parameter at its slice."""

    def __init__(self, allgather_handle, params, partitions: List[torch.Tensor], world_size, use_secondary_tensor=False,
                 quantization=None):
        self.allgather_handle, self.params, self.partitions, self.world_size = allgather_handle, list(params), \
            partitions, world_size
        self.use_secondary_tensor, self.quantization = use_secondary_tensor, quantization
        self.complete = False
        for p in self.params:
            if getattr(p, "ds_status", ZeroParamStatus.INFLIGHT) not in (ZeroParamStatus.INFLIGHT, "INFLIGHT"):
                raise RuntimeError(f"expected param {getattr(p, 'ds_id', '?')} to not be available")

    def wait(self, handle_dependency=True):
        if self.complete:
            return
        if self.allgather_handle is not None and hasattr(self.allgather_handle, "wait"):
            self.allgather_handle.wait()
        offset = 0
        for p in self.params:
            per_rank = getattr(getattr(p, "ds_tensor", None), "ds_numel", None)
            numel = getattr(p, "ds_numel", None)
            if self.partitions and per_rank is not None and numel is not None:
                pieces = [part.narrow(0, offset, per_rank) for part in self.partitions]
                p.data = torch.cat(pieces)[:numel].view(getattr(p, "ds_shape", (numel, ))).to(p.device)
                offset += per_rank
            p.ds_status = ZeroParamStatus.AVAILABLE
        self.complete = True


class MultipleAllGatherHandles:

    def __init__(self, handles):
        self.handles = list(handles)

    def wait(self, handle_dependency=True):
        for h in self.handles:
            h.wait(handle_dependency)


class AllReduceCoalescedHandle:
    """Reduction of several tensors through one flat buffer: ``wait()`` copies the reduced values back."""

    def __init__(self, handle, params):
        self.handle, self.params, self.complete = handle, list(params), False

    def wait(self, **kwargs):
        if self.complete:
            return
        inst = self.handle
        if inst is not None and hasattr(inst, "wait"):
            inst.wait()
        self.complete = True


@dataclass
class QuantizationInfo:
    """What a quantised gather needs to turn its payload back into the parameter."""
    quantized_param: Any = None
    backend: Any = None
    quant_handle: Any = None
    scale_buffer: Any = None
    partition_sz: int = 0
    world_size: int = 1


class ZeRoTraceMode(Enum):
    RECORD = 1  # the coordinator is recording which units a step touches, in order
    COMPLETE = 2  # the recorded order is used to prefetch
    INVALID = 3  # the step deviated from the record: prefetching is off until a new record exists


class InflightParamRegistry(OrderedDict):
    """Parameters whose gather has been launched, with the handle to wait on."""

    def __setitem__(self, param, handle):
        if param in self:
            raise RuntimeError(f"{getattr(param, 'ds_id', param)} already in registry")
        if getattr(param, "ds_status", ZeroParamStatus.INFLIGHT) not in (ZeroParamStatus.INFLIGHT, "INFLIGHT"):
            raise RuntimeError(f"attempted to add non-inflight parameter to registry {getattr(param, 'ds_id', param)}")
        super().__setitem__(param, handle)


class ZeROOrderedDict(OrderedDict):
    """``module._parameters`` replacement that brings a partitioned parameter in when it is read outside the hooks'
    reach (a parent module touching a child's weight directly). Reference ``runtime/zero/parameter_offload.py:29``."""

    def __init__(self, parent_module, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._parent_module = parent_module
        self._in_forward = False

    def __reduce__(self):
        r0, _, *rest = super().__reduce__()
        return (r0, (self._parent_module, ), *rest)

    def __getitem__(self, key):
        param = super().__getitem__(key)
        if param is None or not hasattr(param, "ds_status"):
            return param
        if getattr(self._parent_module, "_parameters", None) is not None and self._in_forward and \
                param.ds_status in (ZeroParamStatus.NOT_AVAILABLE, "NOT_AVAILABLE") and hasattr(param, "all_gather"):
            param.all_gather()
        return param
