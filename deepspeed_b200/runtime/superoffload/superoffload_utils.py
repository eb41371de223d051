"""SuperOffload: speculative CPU optimizer steps overlapped with backward, with rollback.

Reference: ``runtime/superoffload/superoffload_stage3.py:27`` + ``superoffload_utils.py:38`` (a worker *process* runs
CPU Adam per sub-group while the GPU continues the backward pass; if the global gradient norm later demands clipping
or an overflow is found, the speculative updates are undone with ``ds_adam_rollback`` and redone).

Here the worker is a *thread*: the native CPU optimizer (``csrc/cpu/cpu_optim.cpp``) runs without the GIL, host memory
is shared, and no tensors have to cross a process boundary. One task per unit: wait for the unit's gradient D2H copy
(CUDA event), then step the unit's segments."""
import queue
import threading

import torch


class SuperOffloadCPUOptimizer:

    def __init__(self, cpu_optimizer, cpuadam_cores_perc=0.8):
        self.opt = cpu_optimizer
        assert hasattr(cpu_optimizer, "rollback"), "SuperOffload needs DeepSpeedCPUAdam (rollback support)"
        self.tasks = queue.Queue()
        self.done = queue.Queue()
        self.pending = 0
        self.speculated = []  # [(grads dict, lp dict, first)] of this iteration, for rollback
        self._error = None
        self.thread = threading.Thread(target=self._loop, name="superoffload-cpu-adam", daemon=True)
        self.thread.start()

    def _loop(self):
        while True:
            item = self.tasks.get()
            if item is None:
                return
            event, grads, lp, scale, first = item
            try:
                if event is not None:
                    event.synchronize()
                self.opt.step_flat(grads=grads, lp_params=lp, grad_scale=scale, advance_step=first)
            except Exception as e:  # noqa: BLE001
                self._error = e
            self.done.put(1)

    def async_step(self, grads, lp, event=None, grad_scale=1.0):
        """Queue the optimizer update of one unit (its gradient copy is finished when ``event`` fires)."""
        first = len(self.speculated) == 0
        self.speculated.append((grads, lp))
        self.tasks.put((event, grads, lp, grad_scale, first))
        self.pending += 1

    def wait(self):
        while self.pending:
            self.done.get()
            self.pending -= 1
        if self._error is not None:
            err, self._error = self._error, None
            raise err

    def commit(self):
        """The speculation was right (no clipping, no overflow)."""
        self.wait()
        self.speculated = []

    def rollback_and_redo(self, grad_scale, skip=False):
        """Undo every speculative update of this iteration; unless ``skip`` (overflow) redo them with ``grad_scale``."""
        self.wait()
        merged = {}
        for grads, _ in self.speculated:
            merged.update(grads)
        if merged:
            self.opt.rollback(merged, 1.0)
        if not skip:
            first = True
            for grads, lp in self.speculated:
                self.opt.step_flat(grads=grads, lp_params=lp, grad_scale=grad_scale, advance_step=first)
                first = False
        self.speculated = []

    def close(self):
        self.tasks.put(None)


class TaskKeys:
    """Field names of a work item handed to the CPU optimizer worker (reference ``superoffload_utils.py:19``)."""
    PARAM_DATA = "param_data"
    PARAM_GRAD = "param_grad"
    PARAM_GROUP_ID = "param_group_id"
    SUB_GROUP_ID = "sub_group_id"
    ROLLBACK = "rollback"
    LR = "lr"


class ResultKeys:
    UPDATED_PARAM = "updated_param"
    EVENT_TYPE = "event_type"


class EventTypes:
    ADAM_STEP = "adam_step"
    ROLLBACK = "rollback"
