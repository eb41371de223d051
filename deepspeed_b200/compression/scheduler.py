"""Turns each compression technique on once training reaches its ``schedule_offset`` (reference
``compression/scheduler.py:12``). A table maps technique -> the module flag it flips, instead of one method each."""
from ..utils.logging import logger
from .compress import get_module_name
from .config import get_compression_config
from .constants import *  # noqa: F401,F403
from .helper import recursive_getattr

_FLAG = {
    WEIGHT_QUANTIZATION: "weight_quantization_enabled",
    ACTIVATION_QUANTIZATION: "activation_quantization_enabled",
    SPARSE_PRUNING: "sparse_pruning_enabled",
    HEAD_PRUNING: "head_pruning_enabled",
    ROW_PRUNING: "row_pruning_enabled",
    CHANNEL_PRUNING: "channel_pruning_enabled",
}


class compression_scheduler:

    def __init__(self, model, compression_config):
        self.model = model
        if SHARED_PARAMETERS not in (compression_config.get(WEIGHT_QUANTIZATION) or {}):
            compression_config = get_compression_config({COMPRESSION_TRAINING: compression_config})
        self.compression_config = compression_config
        self.make_init()
        self.training_steps = 0
        self.weight_quantization_enabled = False
        self.verbose = {t: False for t in _FLAG}

    def make_init(self):
        self.different_compression_methods = {}
        for method, content in self.compression_config.items():
            if LAYER_REDUCTION in method:
                continue
            shared = content[SHARED_PARAMETERS]
            entry = {TECHNIQUE_ENABLED: shared[TECHNIQUE_ENABLED], SHARED_PARAMETERS: shared, DIFFERENT_GROUPS: []}
            seen = set()
            for group_name, gp in content[DIFFERENT_GROUPS].items():
                names = []
                for key_word in gp[DIFFERENT_GROUPS_MODULE_SCOPE]:
                    found, seen = get_module_name(group_name, self.model, key_word, seen, verbose=False)
                    names.extend(found)
                if names:
                    entry[DIFFERENT_GROUPS].append([group_name, names, dict(gp[DIFFERENT_GROUPS_PARAMETERS])])
            self.different_compression_methods[method] = entry

    def _due(self, method, shared):
        start = shared[TECHNIQUE_SCHEDULE_OFFSET]
        if method == SPARSE_PRUNING:
            end = shared.get(TECHNIQUE_SCHEDULE_OFFSET_END, start)
            if shared.get(SPARSE_PRUNING_METHOD) != SPARSE_PRUNING_METHOD_SNIP_MOMENTUM:
                return self.training_steps >= start
            return start <= self.training_steps <= end
        return self.training_steps >= start

    def check(self, method):
        entry = self.different_compression_methods[method]
        if not entry[TECHNIQUE_ENABLED] or not self._due(method, entry[SHARED_PARAMETERS]):
            return
        for _, names, _ in entry[DIFFERENT_GROUPS]:
            for name in names:
                setattr(recursive_getattr(self.model, name), _FLAG[method], True)
        if not self.verbose[method]:
            logger.info(f"{method} is enabled at step {self.training_steps}")
            self.verbose[method] = True
            if method == WEIGHT_QUANTIZATION:
                self.weight_quantization_enabled = True

    def check_weight_quantization(self):
        self.check(WEIGHT_QUANTIZATION)

    def check_activation_quantization(self):
        self.check(ACTIVATION_QUANTIZATION)

    def check_sparse_pruning(self):
        self.check(SPARSE_PRUNING)

    def check_head_pruning(self):
        self.check(HEAD_PRUNING)

    def check_row_pruning(self):
        self.check(ROW_PRUNING)

    def check_channel_pruning(self):
        self.check(CHANNEL_PRUNING)

    def check_all_modules(self):
        for method in _FLAG:
            self.check(method)

    def step(self, step_zero_check=False):
        if not step_zero_check:
            self.training_steps += 1
        self.check_all_modules()
