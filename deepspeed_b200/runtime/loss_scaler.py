"""Static and dynamic loss scaling for fp16.

Semantics follow the reference ``runtime/fp16/loss_scaler.py`` (``LossScaler`` static, ``DynamicLossScaler``
:187 with scale window / hysteresis (``delayed_shift``) / ``consecutive_hysteresis`` / ``min_scale``,
``CreateLossScaler`` :304). Objects are picklable: they are stored inside optimizer checkpoints.
"""
import torch

INITIAL_LOSS_SCALE = "init_scale"
SCALE_WINDOW = "scale_window"
DELAYED_SHIFT = "delayed_shift"
CONSECUTIVE_HYSTERESIS = "consecutive_hysteresis"
MIN_LOSS_SCALE = "min_scale"


class LossScalerBase:

    def __init__(self, cur_scale):
        self.cur_scale = cur_scale
        self.dynamic = False

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        return tuple(self.loss_scale * g for g in grad_in)

    def update_scale(self, overflow):
        pass

    def backward(self, loss, retain_graph=False):
        (loss * self.loss_scale).backward(retain_graph=retain_graph)


class LossScaler(LossScalerBase):
    """Static loss scale."""

    def __init__(self, scale=1):
        super().__init__(scale)

    def has_overflow(self, params):
        return False

    @staticmethod
    def _has_inf_or_nan(x):
        return False


class DynamicLossScaler(LossScalerBase):
    """Halve on overflow (after ``delayed_shift`` hysteresis credits are used up), double after
    ``scale_window`` clean steps."""

    def __init__(self, init_scale=2**32, scale_factor=2.0, scale_window=1000, min_scale=1, delayed_shift=1,
                 consecutive_hysteresis=False, raise_error_at_min_scale=True, dtype=torch.half):
        super().__init__(init_scale)
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = delayed_shift
        self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale
        self.dynamic = True
        self.dtype = dtype

    @staticmethod
    def _has_inf_or_nan(x):
        try:
            s = float(x.float().sum())
        except RuntimeError as e:
            if "value cannot be converted" not in e.args[0]:
                raise
            return True
        return s in (float("inf"), -float("inf")) or s != s

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    def update_scale(self, overflow):
        if overflow:
            if self.delayed_shift == 1 or self.cur_hysteresis == 1:
                if self.cur_scale == self.min_scale and self.raise_error_at_min_scale:
                    raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. "
                                    "Exiting run.")
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_scale)
            else:
                self.cur_hysteresis -= 1
            self.last_overflow_iter = self.cur_iter
        else:
            if self.consecutive_hysteresis:
                self.cur_hysteresis = self.delayed_shift
            if (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0:
                if not self.consecutive_hysteresis:
                    self.cur_hysteresis = self.delayed_shift
                self.cur_scale *= self.scale_factor
        self.cur_iter += 1


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    if dtype == torch.half and dynamic_scaling:
        assert dynamic_loss_args is not None, "Dynamic loss scaling parameters must be defined."
        return DynamicLossScaler(dtype=dtype, **dynamic_loss_args)
    loss_scale_value = static_loss_scale if dtype == torch.half else 1.0
    return LossScaler(scale=loss_scale_value)


# Pickle these small objects under the reference's module path so optimizer checkpoints are interchangeable with the
# reference's tools/engine (``deepspeed.runtime.fp16.loss_scaler``; resolved here by the ``deepspeed`` alias package).
for _cls in (LossScalerBase, LossScaler, DynamicLossScaler):
    _cls.__module__ = "deepspeed.runtime.fp16.loss_scaler"


# ---- loss-scale bookkeeping of the fp16 optimizer wrappers (reference ``fp16/loss_scaler.py:39-122``) ---------
from dataclasses import dataclass  # noqa: E402
from enum import Enum  # noqa: E402


class LossScaleProfile(str, Enum):
    FUSED = "fused"  # FP16_Optimizer: starts at 2^32
    UNFUSED = "unfused"  # FP16_UnfusedOptimizer: starts at 2^16, may go below 1


@dataclass(frozen=True)
class LossScaleProfileDefaults:
    initial_dynamic_scale: float
    default_scale_window: int
    default_min_loss_scale: float
    scale_factor: float


LOSS_SCALE_PROFILE_DEFAULTS = {
    LossScaleProfile.FUSED: LossScaleProfileDefaults(2**32, 1000, 1, 2.0),
    LossScaleProfile.UNFUSED: LossScaleProfileDefaults(1.0 * 2**16, 1000, 0.25, 2.0),
}


class LossScaleConfig:
    """Resolved loss-scaling state: no scaling unless the low-precision dtype is fp16; static scale, or dynamic with
    the profile's defaults / the user's ``dynamic_loss_args`` (``init_scale``, ``scale_window``, ``min_scale``)."""

    def __init__(self, low_precision_dtype, dynamic_loss_scale, static_loss_scale, dynamic_loss_args, *,
                 profile=LossScaleProfile.FUSED, initial_dynamic_scale=None):
        import torch
        d = LOSS_SCALE_PROFILE_DEFAULTS[profile]
        self.use_grad_scaling = low_precision_dtype == torch.float16
        self.dynamic_loss_scale = False
        self.cur_iter, self.cur_scale = 0, 1.0
        self.last_overflow_iter = self.scale_factor = self.scale_window = self.min_loss_scale = None
        if not self.use_grad_scaling:
            return
        self.cur_scale = static_loss_scale
        if not dynamic_loss_scale:
            return
        self.dynamic_loss_scale = True
        self.last_overflow_iter = -1
        self.scale_factor = d.scale_factor
        if dynamic_loss_args is None:
            self.cur_scale = d.initial_dynamic_scale if initial_dynamic_scale is None else initial_dynamic_scale
            self.scale_window, self.min_loss_scale = d.default_scale_window, d.default_min_loss_scale
        else:
            self.cur_scale = dynamic_loss_args[INITIAL_LOSS_SCALE]
            self.scale_window = dynamic_loss_args[SCALE_WINDOW]
            self.min_loss_scale = dynamic_loss_args[MIN_LOSS_SCALE]


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t[0]
