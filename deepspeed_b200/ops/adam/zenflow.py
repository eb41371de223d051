"""Stand-alone ZenFlow optimizers (reference ``ops/adam/zenflow_cpu_adam.py``, ``zenflow_torch_adam.py``). Inside the
engine ZenFlow runs through ``runtime/zenflow/zenflow_utils.py::ZenFlowController`` on the flat shards; these classes give
user code the reference's per-parameter form: AdamW applied only to the *selected* columns of each 2-D weight
(``param.selected_indices`` + ``param.selected_grad``; 1-D parameters are updated whole), with the moments of the
selection optionally parked on the host between steps."""
import threading

import torch

from ..cpu_optimizers import DeepSpeedCPUAdam


class ZenFlowCPUAdam(DeepSpeedCPUAdam):
    """CPU Adam whose step can run while the GPU already works on the next iteration (``overlap_step``): the step is
    handed to a worker thread and joined at the next call / at ``wait()``."""

    def __init__(self, *args, overlap_step=False, **kwargs):
        super().__init__(*args, **kwargs)
        self.overlap_step = overlap_step
        self._worker = None

    def wait(self):
        if self._worker is not None:
            self._worker.join()
            self._worker = None

    def step(self, step_id=None, closure=None, **kwargs):
        self.wait()
        if not self.overlap_step:
            return super().step(closure) if closure is not None else super().step()
        self._worker = threading.Thread(target=lambda: DeepSpeedCPUAdam.step(self), daemon=True)
        self._worker.start()
        return None


class ZenFlowSelectiveAdamW(torch.optim.AdamW):

    def __init__(self, *args, offload=False, bucket_size=5e8, **kwargs):
        super().__init__(*args, **kwargs)
        self.offload = offload
        self.bucket_size = int(bucket_size)

    @staticmethod
    def _selected_view(param):
        if param.dim() == 1 or not hasattr(param, "selected_indices"):
            return param.data, None
        return param.data[:, param.selected_indices], param.selected_indices

    def temp_copy_param(self, group_to_paramlist):
        """Snapshot the selected columns (the CPU optimizer is about to overwrite the whole parameter)."""
        for params in group_to_paramlist.values():
            for p in params:
                if hasattr(p, "selected_grad"):
                    snap = self._selected_view(p)[0].clone().detach()
                    p.temp_selected_param = snap.cpu() if self.offload else snap

    def clear_selected_mv(self):
        for group in self.param_groups:
            for p in group["params"]:
                st = self.state.get(p)
                if st:
                    st["exp_avg"].zero_()
                    st["exp_avg_sq"].zero_()

    @torch.no_grad()
    def step(self, closure=None):
        for group in self.param_groups:
            b1, b2 = group["betas"]
            lr, wd, eps = group["lr"], group["weight_decay"], group["eps"]
            for p in group["params"]:
                g = getattr(p, "selected_grad", None)
                if g is None:
                    continue
                sel, idx = self._selected_view(p)
                if getattr(p, "temp_selected_param", None) is not None:
                    sel = p.temp_selected_param.to(p.device)
                st = self.state.setdefault(p, {})
                if not st:
                    home = "cpu" if self.offload else sel.device
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(sel, device=home)
                    st["exp_avg_sq"] = torch.zeros_like(sel, device=home)
                m, v = st["exp_avg"].to(sel.device), st["exp_avg_sq"].to(sel.device)
                st["step"] += 1
                g = g.to(sel.dtype)
                m.mul_(b1).add_(g, alpha=1 - b1)
                v.mul_(b2).addcmul_(g, g, value=1 - b2)
                bc1, bc2 = 1 - b1**st["step"], 1 - b2**st["step"]
                new = sel * (1 - lr * wd) - lr * (m / bc1) / ((v / bc2).sqrt() + eps)
                if idx is None:
                    p.data.copy_(new)
                else:
                    p.data[:, idx] = new
                st["exp_avg"].copy_(m)
                st["exp_avg_sq"].copy_(v)
                p.temp_selected_param = None
                p.selected_grad = None

    def group_step(self, group_to_paramlist):
        """Step only the listed parameters (called bucket by bucket as their gradients are reduced)."""
        wanted = {id(p) for ps in group_to_paramlist.values() for p in ps}
        held = []
        for group in self.param_groups:
            for p in group["params"]:
                if id(p) not in wanted and getattr(p, "selected_grad", None) is not None:
                    held.append((p, p.selected_grad))
                    p.selected_grad = None
        self.step()
        for p, g in held:
            p.selected_grad = g


class ZenFlowSelectiveAdamW_stage3(ZenFlowSelectiveAdamW):
    """ZeRO-3 form: parameters are flat partitions, the selection indexes elements (``param.selected_indices`` over the
    flattened partition) instead of columns."""

    @staticmethod
    def _selected_view(param):
        if not hasattr(param, "selected_indices"):
            return param.data, None
        return param.data.view(-1)[param.selected_indices], param.selected_indices

    def temp_copy_param(self, paramlist):
        super().temp_copy_param({0: list(paramlist)})

    def group_step(self, paramlist):
        super().group_step({0: list(paramlist)})

    @torch.no_grad()
    def step(self, closure=None):
        flat_idx = {}
        for group in self.param_groups:
            for p in group["params"]:
                if getattr(p, "selected_grad", None) is not None and hasattr(p, "selected_indices") and p.dim() != 1:
                    flat_idx[p] = (p.data, p.selected_indices)
        # run the column form on flattened views
        for p, (data, idx) in flat_idx.items():
            p.data = data.view(1, -1)
        try:
            super().step(closure)
        finally:
            for p, (data, idx) in flat_idx.items():
                p.data = data
