"""``OnDevice``: construct a model directly on a device / in a dtype, or as meta tensors (reference
``utils/init_on_device.py``). Implemented with torch's own device context plus a default-dtype switch instead of
patching ``torch.empty`` and friends."""
import torch


class OnDevice:
    """``with OnDevice(dtype=torch.bfloat16, device="meta"): model = MyModel()``"""

    def __init__(self, dtype, device="meta", enabled=True):
        self.dtype, self.device, self.enabled = dtype, device, enabled
        self._dev_ctx = None
        self._old_dtype = None

    def __enter__(self):
        if not self.enabled:
            return self
        self._dev_ctx = torch.device(self.device)
        self._dev_ctx.__enter__()
        if self.dtype is not None and self.dtype.is_floating_point:
            self._old_dtype = torch.get_default_dtype()
            torch.set_default_dtype(self.dtype)
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        if not self.enabled:
            return False
        if self._old_dtype is not None:
            torch.set_default_dtype(self._old_dtype)
            self._old_dtype = None
        self._dev_ctx.__exit__(exc_type, exc_value, traceback)
        self._dev_ctx = None
        return False
