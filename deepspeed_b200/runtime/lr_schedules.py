"""Learning-rate schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR, WarmupCosineLR.

Parameter names and curve shapes follow the reference ``runtime/lr_schedules.py`` (classes at :288, :389,
:678, :766, :817; argparse helpers ``add_tuning_arguments``). Implemented on one small base class.
"""
import argparse
import math

from ..utils.logging import logger

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST = "LRRangeTest"
ONE_CYCLE = "OneCycle"
WARMUP_LR = "WarmupLR"
WARMUP_DECAY_LR = "WarmupDecayLR"
WARMUP_COSINE_LR = "WarmupCosineLR"
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR, WARMUP_COSINE_LR]

LR_RANGE_TEST_MIN_LR = "lr_range_test_min_lr"
LR_RANGE_TEST_STEP_RATE = "lr_range_test_step_rate"
LR_RANGE_TEST_STEP_SIZE = "lr_range_test_step_size"
LR_RANGE_TEST_STAIRCASE = "lr_range_test_staircase"
CYCLE_MIN_LR = "cycle_min_lr"
CYCLE_MAX_LR = "cycle_max_lr"
WARMUP_MIN_LR = "warmup_min_lr"
WARMUP_MAX_LR = "warmup_max_lr"
WARMUP_NUM_STEPS = "warmup_num_steps"
WARMUP_TYPE = "warmup_type"
WARMUP_LOG_RATE = "log"
WARMUP_LINEAR_RATE = "linear"
TOTAL_NUM_STEPS = "total_num_steps"


def add_tuning_arguments(parser):
    group = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    group.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    group.add_argument("--lr_range_test_min_lr", type=float, default=0.001)
    group.add_argument("--lr_range_test_step_rate", type=float, default=1.0)
    group.add_argument("--lr_range_test_step_size", type=int, default=1000)
    group.add_argument("--lr_range_test_staircase", type=bool, default=False)
    group.add_argument("--cycle_first_step_size", type=int, default=1000)
    group.add_argument("--cycle_first_stair_count", type=int, default=-1)
    group.add_argument("--cycle_second_step_size", type=int, default=-1)
    group.add_argument("--cycle_second_stair_count", type=int, default=-1)
    group.add_argument("--decay_step_size", type=int, default=1000)
    group.add_argument("--cycle_min_lr", type=float, default=0.01)
    group.add_argument("--cycle_max_lr", type=float, default=0.1)
    group.add_argument("--decay_lr_rate", type=float, default=0.0)
    group.add_argument("--cycle_momentum", default=False, action="store_true")
    group.add_argument("--cycle_min_mom", type=float, default=0.8)
    group.add_argument("--cycle_max_mom", type=float, default=0.9)
    group.add_argument("--decay_mom_rate", type=float, default=0.0)
    group.add_argument("--warmup_min_lr", type=float, default=0)
    group.add_argument("--warmup_max_lr", type=float, default=0.001)
    group.add_argument("--warmup_num_steps", type=int, default=1000)
    group.add_argument("--warmup_type", type=str, default=WARMUP_LOG_RATE)
    return parser


def parse_arguments():
    parser = argparse.ArgumentParser()
    parser = add_tuning_arguments(parser)
    return parser.parse_known_args()


def get_torch_optimizer(optimizer):
    import torch
    if isinstance(optimizer, torch.optim.Optimizer):
        return optimizer
    if hasattr(optimizer, "optimizer") and isinstance(optimizer.optimizer, torch.optim.Optimizer):
        return optimizer.optimizer
    if hasattr(optimizer, "param_groups"):
        return optimizer
    raise TypeError(f"{type(optimizer).__name__} is not a subclass of torch.optim.Optimizer")


def _as_list(optimizer, name, value):
    n = len(optimizer.param_groups)
    if isinstance(value, (list, tuple)):
        if len(value) != n:
            raise ValueError(f"Expected {n} values for {name}, got {len(value)}")
        return list(value)
    return [value] * n


class _Schedule:
    """Common plumbing: ``step(last_batch_iteration=None)`` writes ``get_lr()`` into the param groups."""

    def __init__(self, optimizer, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.last_batch_iteration = last_batch_iteration
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def get_lr(self):
        raise NotImplementedError

    def get_last_lr(self):
        assert getattr(self, "_last_lr", None) is not None, "need to call step() first"
        return self._last_lr

    def _apply(self):
        lrs = self.get_lr()
        for g, lr in zip(self.optimizer.param_groups, lrs):
            g["lr"] = lr
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        self._apply()

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]


class LRRangeTest(_Schedule):
    """lr = min_lr * (1 + step_rate * interval(t)); interval continuous or staircase."""

    def __init__(self, optimizer, lr_range_test_min_lr=1e-3, lr_range_test_step_size=2000,
                 lr_range_test_step_rate=1.0, lr_range_test_staircase=False, last_batch_iteration=-1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lr = _as_list(self.optimizer, "lr_range_test_min_lr", lr_range_test_min_lr)
        self.step_size = lr_range_test_step_size
        self.step_rate = lr_range_test_step_rate
        self.staircase = lr_range_test_staircase
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lr):
                g["lr"] = lr
            self._last_lr = list(self.min_lr)

    def _interval(self):
        x = float(self.last_batch_iteration + 1) / self.step_size
        return math.floor(x) if self.staircase else x

    def get_lr(self):
        inc = 1 + self.step_rate * self._interval()
        return [lr * inc for lr in self.min_lr]


class OneCycle(_Schedule):
    """1-cycle policy: lr min->max over first phase, max->min over second, then decay; inverse for momentum."""

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate=0.0, cycle_first_step_size=2000,
                 cycle_second_step_size=None, cycle_first_stair_count=0, cycle_second_stair_count=None,
                 decay_step_size=0, cycle_momentum=True, cycle_min_mom=0.8, cycle_max_mom=0.9, decay_mom_rate=0.0,
                 last_batch_iteration=-1):
        super().__init__(optimizer, last_batch_iteration)
        first = float(cycle_first_step_size)
        second = float(cycle_second_step_size) if cycle_second_step_size is not None else first
        self.total_size = first + second
        self.step_ratio = first / self.total_size
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None \
            else cycle_second_stair_count
        self.decay_step_size = decay_step_size
        n = len(self.optimizer.param_groups)
        self.min_lrs, self.max_lrs = [cycle_min_lr] * n, [cycle_max_lr] * n
        self.decay_lr_rate = decay_lr_rate
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lrs):
                g["lr"] = lr
        self.cycle_momentum = cycle_momentum
        if cycle_momentum:
            key = "betas" if "betas" in self.optimizer.param_groups[0] else (
                "momentum" if "momentum" in self.optimizer.param_groups[0] else None)
            if key is None:
                logger.warning(f"cycle_momentum is disabled because optimizer {type(self.optimizer).__name__} does "
                               f"not support momentum, no betas attribute in defaults")
                self.cycle_momentum = False
            else:
                self._mom_key = key
                self.decay_mom_rate = decay_mom_rate
                self.min_moms = [(cycle_min_mom, 0.99)] * n
                self.max_moms = [(cycle_max_mom, 0.99)] * n
                if last_batch_iteration == -1:
                    for g in self.optimizer.param_groups:
                        self._set_mom(g, cycle_max_mom)
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]

    def _set_mom(self, group, value):
        if self._mom_key == "betas":
            group["betas"] = (value, group["betas"][1])
        else:
            group["momentum"] = value

    def _get_scale_factor(self):
        it = self.last_batch_iteration + 1
        cycle = math.floor(1 + it / self.total_size)
        x = 1.0 + it / self.total_size - cycle
        return x / self.step_ratio if x <= self.step_ratio else (x - 1) / (self.step_ratio - 1)

    def _cycle_lr(self):
        sf = self._get_scale_factor()
        return [mn + (mx - mn) * sf for mn, mx in zip(self.min_lrs, self.max_lrs)]

    def _cycle_mom(self):
        sf = self._get_scale_factor()
        return [mx[0] - (mx[0] - mn[0]) * sf for mn, mx in zip(self.min_moms, self.max_moms)]

    def _decay_interval(self):
        decay_it = self.last_batch_iteration - self.total_size + 1
        return decay_it / self.decay_step_size if self.decay_step_size else 0.0

    def get_lr(self):
        if self.last_batch_iteration < self.total_size:
            return self._cycle_lr()
        f = 1 + self.decay_lr_rate * self._decay_interval()
        return [mn / f for mn in self.min_lrs]

    def get_mom(self):
        if not self.cycle_momentum:
            return None
        if self.last_batch_iteration < self.total_size:
            return self._cycle_mom()
        f = 1 + self.decay_mom_rate * self._decay_interval()
        return [mx[0] * f for mx in self.max_moms]

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply()
        if self.cycle_momentum:
            for g, m in zip(self.optimizer.param_groups, self.get_mom()):
                self._set_mom(g, m)


class WarmupLR(_Schedule):
    """min_lr -> max_lr over ``warmup_num_steps`` (log or linear), then constant."""

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        super().__init__(optimizer, last_batch_iteration)
        self.min_lrs = _as_list(self.optimizer, "min_lr", warmup_min_lr)
        self.max_lrs = _as_list(self.optimizer, "max_lr", warmup_max_lr)
        self.delta_lrs = [big - small for big, small in zip(self.max_lrs, self.min_lrs)]
        if warmup_type not in {WARMUP_LOG_RATE, WARMUP_LINEAR_RATE}:
            logger.warning(f"Using unknown warmup_type: {warmup_type}. The increasing function is set to default (log)")
            warmup_type = WARMUP_LOG_RATE
        self.warmup_type = warmup_type
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if last_batch_iteration == -1:
            self._last_lr = self._apply_at(-1)

    def _apply_at(self, it):
        self.last_batch_iteration = it
        self._apply()
        return self._last_lr

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            return self.last_batch_iteration / self.warmup_num_steps
        return 1.0

    def get_lr(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return list(self.min_lrs)
        gamma = self._get_gamma()
        return [mn + d * gamma for mn, d in zip(self.min_lrs, self.delta_lrs)]


class WarmupDecayLR(WarmupLR):
    """Warmup then linear decay to 0 at ``total_num_steps``."""

    def __init__(self, optimizer, total_num_steps, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, warmup_type, last_batch_iteration)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning(f"total_num_steps {total_num_steps} is less than warmup_num_steps {warmup_num_steps}")

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return super()._get_gamma()
        return max(0.0, float(self.total_num_steps - self.last_batch_iteration) /
                   float(max(1.0, self.total_num_steps - self.warmup_num_steps)))


class WarmupCosineLR(_Schedule):
    """Linear/log warmup of the lr *ratio* from ``warmup_min_ratio`` to 1, then cosine to ``cos_min_ratio``."""

    def __init__(self, optimizer, total_num_steps, warmup_min_ratio=0.0, warmup_num_steps=1000, cos_min_ratio=0.0001,
                 warmup_type=WARMUP_LOG_RATE, last_batch_iteration=-1):
        super().__init__(optimizer, last_batch_iteration)
        self.total_num_steps = total_num_steps
        self.warmup_min_ratio = warmup_min_ratio
        self.cos_min_ratio = cos_min_ratio
        if warmup_type not in {WARMUP_LOG_RATE, WARMUP_LINEAR_RATE}:
            warmup_type = WARMUP_LOG_RATE
        self.warmup_type = warmup_type
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning(f"total_num_steps {total_num_steps} is less than warmup_num_steps {warmup_num_steps}")
        self.org_lrs = [g["lr"] for g in self.optimizer.param_groups]
        if last_batch_iteration == -1:
            self.last_batch_iteration = -1
            self._apply()

    def get_lr_ratio(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return 0.0
        if self.last_batch_iteration < self.warmup_num_steps:
            if self.warmup_type == WARMUP_LOG_RATE:
                r = self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
            else:
                r = self.last_batch_iteration / self.warmup_num_steps
            return self.warmup_min_ratio + (1.0 - self.warmup_min_ratio) * r
        real_last = self.last_batch_iteration - self.warmup_num_steps + 1
        real_total = self.total_num_steps - self.warmup_num_steps
        ratio_delta = 1.0 - self.cos_min_ratio
        ratio = (1 + math.cos(math.pi * real_last / real_total)) / 2
        return max(0.0, self.cos_min_ratio + ratio_delta * ratio)

    def get_lr(self):
        if self.last_batch_iteration < 0:
            return [0.0 for _ in self.org_lrs]
        r = self.get_lr_ratio()
        return [lr * r for lr in self.org_lrs]


_CLASSES = {LR_RANGE_TEST: LRRangeTest, ONE_CYCLE: OneCycle, WARMUP_LR: WarmupLR, WARMUP_DECAY_LR: WarmupDecayLR,
            WARMUP_COSINE_LR: WarmupCosineLR}


def get_lr_schedule_class(name):
    return _CLASSES.get(name)


def build_lr_scheduler(name, optimizer, params):
    """Instantiate a schedule by ds_config name, or a ``torch.optim.lr_scheduler`` class of that name."""
    if name in _CLASSES:
        return _CLASSES[name](optimizer, **params)
    import torch
    if hasattr(torch.optim.lr_scheduler, name):
        return getattr(torch.optim.lr_scheduler, name)(get_torch_optimizer(optimizer), **params)
    raise ValueError(f"Unknown scheduler type {name}; valid: {VALID_LR_SCHEDULES} or torch.optim.lr_scheduler.*")


# ---- command line <-> schedule config (reference ``lr_schedules.py:133-258``) -------------------------------
CYCLE_FIRST_STEP_SIZE = "cycle_first_step_size"
CYCLE_FIRST_STAIR_COUNT = "cycle_first_stair_count"
CYCLE_SECOND_STEP_SIZE = "cycle_second_step_size"
CYCLE_SECOND_STAIR_COUNT = "cycle_second_stair_count"
DECAY_STEP_SIZE = "decay_step_size"
DECAY_LR_RATE = "decay_lr_rate"
CYCLE_MIN_MOM = "cycle_min_mom"
CYCLE_MAX_MOM = "cycle_max_mom"
DECAY_MOM_RATE = "decay_mom_rate"

_LR_RANGE_KEYS = (LR_RANGE_TEST_MIN_LR, LR_RANGE_TEST_STEP_RATE, LR_RANGE_TEST_STEP_SIZE, LR_RANGE_TEST_STAIRCASE)
_ONE_CYCLE_KEYS = (CYCLE_FIRST_STEP_SIZE, CYCLE_FIRST_STAIR_COUNT, CYCLE_SECOND_STEP_SIZE, CYCLE_SECOND_STAIR_COUNT,
                   DECAY_STEP_SIZE, CYCLE_MIN_LR, CYCLE_MAX_LR, DECAY_LR_RATE, CYCLE_MIN_MOM, CYCLE_MAX_MOM,
                   DECAY_MOM_RATE)
_WARMUP_KEYS = (WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS, WARMUP_TYPE)


def _override(args, params, keys):
    for k in keys:
        v = getattr(args, k, None)
        if v is not None:
            params[k] = v


def override_lr_range_test_params(args, params):
    _override(args, params, _LR_RANGE_KEYS)


def override_1cycle_params(args, params):
    _override(args, params, _ONE_CYCLE_KEYS)


def override_warmupLR_params(args, params):
    _override(args, params, _WARMUP_KEYS)


def override_params(args, params):
    """Command-line values (``add_tuning_arguments``) win over the values of a schedule config."""
    override_lr_range_test_params(args, params)
    override_1cycle_params(args, params)
    override_warmupLR_params(args, params)


def get_config_from_args(args):
    """``(scheduler config dict, None)`` built from ``--lr_schedule`` and its parameters, or ``(None, reason)``."""
    name = getattr(args, LR_SCHEDULE, None)
    if name is None:
        return None, f"--{LR_SCHEDULE} not specified on command line"
    if name not in VALID_LR_SCHEDULES:
        return None, f"{name} is not supported LR schedule"
    config = {"type": name, "params": {}}
    if name == LR_RANGE_TEST:
        override_lr_range_test_params(args, config["params"])
    elif name == ONE_CYCLE:
        override_1cycle_params(args, config["params"])
    else:
        override_warmupLR_params(args, config["params"])
    return config, None


def get_lr_from_config(config):
    """The learning rate a schedule config starts the optimizer with: ``(lr, '')`` or ``(None, reason)``."""
    if "type" not in config:
        return None, "LR schedule type not defined in config"
    if "params" not in config:
        return None, "LR schedule params not defined in config"
    name, params = config["type"], config["params"]
    if name not in VALID_LR_SCHEDULES:
        return None, f"{name} is not a valid LR schedule"
    if name == LR_RANGE_TEST:
        return params[LR_RANGE_TEST_MIN_LR], ""
    if name == ONE_CYCLE:
        return params[CYCLE_MAX_LR], ""
    return params[WARMUP_MAX_LR], ""


def update_lr(param_groups, lrs):
    """Write ``lrs`` into the groups (in place for tensor learning rates, as capturable optimizers use)."""
    import torch
    for group, lr in zip(param_groups, lrs):
        if torch.is_tensor(group["lr"]):
            group["lr"].fill_(lr.squeeze() if torch.is_tensor(lr) else lr)
        else:
            group["lr"] = lr
    return [g["lr"] for g in param_groups]
