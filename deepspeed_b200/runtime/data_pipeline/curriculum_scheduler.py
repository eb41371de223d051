"""Curriculum difficulty schedules (reference ``runtime/data_pipeline/curriculum_scheduler.py:11``):
``fixed_discrete`` (step table), ``fixed_root`` / ``fixed_linear`` (root-of-progress, quantised to ``difficulty_step``)
and ``custom`` (user callback)."""
import math

from ...utils.logging import logger

_REQUIRED = {
    "fixed_discrete": ("difficulty", "max_step"),
    "fixed_root": ("total_curriculum_step", "difficulty_step", "root_degree"),
    "fixed_linear": ("total_curriculum_step", "difficulty_step"),
    "custom": (),
}


class CurriculumScheduler:

    def __init__(self, config):
        for key in ("min_difficulty", "max_difficulty", "schedule_type"):
            assert key in config, f"Curriculum learning requires the config '{key}'"
        kind = config["schedule_type"]
        if kind not in _REQUIRED:
            raise RuntimeError("Unsupported curriculum schedule type")
        self.state = {
            "min_difficulty": config["min_difficulty"],
            "max_difficulty": config["max_difficulty"],
            "current_difficulty": config["min_difficulty"],
            "schedule_type": kind,
        }
        self.first_step = True
        self.custom_get_difficulty = None
        if kind != "custom":
            sc = config["schedule_config"]
            for key in _REQUIRED[kind]:
                assert key in sc, f"Curriculum learning with {kind} schedule requires the schedule_config '{key}'"
            if kind == "fixed_discrete":
                assert len(sc["max_step"]) > 0 and len(sc["difficulty"]) == len(sc["max_step"]) + 1
            elif sc["difficulty_step"] % 8 != 0:
                logger.warning("curriculum difficulty_step is not a multiple of 8: sequence-length curricula then "
                               "produce GEMM shapes the tensor cores handle poorly")
            self.state["schedule_config"] = sc

    def get_current_difficulty(self):
        return self.state["current_difficulty"]

    def set_current_difficulty(self, difficulty):
        self.state["current_difficulty"] = difficulty

    def set_custom_get_difficulty(self, schedule_function):
        self.custom_get_difficulty = schedule_function

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def _discrete(self, step):
        sc = self.state["schedule_config"]
        for limit, diff in zip(sc["max_step"], sc["difficulty"]):
            if step <= limit:
                return diff
        return sc["difficulty"][-1]

    def _root(self, step, degree=None):
        sc = self.state["schedule_config"]
        degree = sc["root_degree"] if degree is None else degree
        lo, hi = self.state["min_difficulty"], self.state["max_difficulty"]
        frac = (float(step) / sc["total_curriculum_step"])**(1.0 / degree)
        nxt = math.floor(frac * (hi - lo) + lo)
        nxt -= nxt % sc["difficulty_step"]
        return min(nxt, hi)

    def get_difficulty(self, global_steps):
        kind = self.state["schedule_type"]
        if kind == "fixed_discrete":
            return self._discrete(global_steps)
        if kind == "fixed_linear":
            return self._root(global_steps, 1)
        if kind == "fixed_root":
            return self._root(global_steps)
        if kind == "custom":
            return self.custom_get_difficulty(global_steps)
        raise RuntimeError("Unsupported curriculum schedule type")

    def update_difficulty(self, global_steps):
        if self.state["current_difficulty"] < self.state["max_difficulty"]:
            self.state["current_difficulty"] = self.get_difficulty(global_steps)
        return self.state["current_difficulty"]
