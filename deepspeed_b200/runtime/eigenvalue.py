"""Power-iteration estimate of the largest Hessian eigenvalue per layer block (reference ``runtime/eigenvalue.py:13``),
used by MoQ to schedule quantization precision per layer."""
import torch

from ..utils.logging import log_dist


class Eigenvalue:

    def __init__(self, verbose=False, max_iter=100, tol=1e-2, stability=0, gas_boundary_resolution=1, layer_name="",
                 layer_num=0):
        self.verbose, self.max_iter, self.tol, self.stability = verbose, max_iter, tol, stability
        self.gas_boundary_resolution, self.layer_name, self.layer_num = gas_boundary_resolution, layer_name, layer_num
        assert len(self.layer_name) > 0 and layer_num > 0
        log_dist(f"enabled eigenvalue with verbose={verbose}, max_iter={max_iter}, tol={tol}, stability={stability}, "
                 f"gas_boundary_resolution={gas_boundary_resolution}, layer_name={layer_name}, layer_num={layer_num}", [0])

    @staticmethod
    def nan_to_num(x):
        return torch.nan_to_num(x, nan=0.0, posinf=0.0, neginf=0.0)

    def normalize(self, v):
        norm = torch.sqrt(sum(torch.sum(x * x) for x in v)) + self.stability
        return [self.nan_to_num(x / norm) for x in v]

    @staticmethod
    def inner_product(xs, ys):
        return sum(torch.sum(x * y) for x, y in zip(xs, ys))

    def get_layers(self, module):
        scope = module
        for name in self.layer_name.split("."):
            scope = getattr(scope, name)
        return scope

    def compute_eigenvalue(self, module, device=None, scale=1.0):
        """Needs a retained graph: call after ``loss.backward(create_graph=True)``. Returns {param_id: (ev, layer)}."""
        block_eigenvalue = []
        layers = self.get_layers(module)
        for block in range(self.layer_num):
            params = [p for p in layers[block].parameters() if p.grad is not None and p.grad.grad_fn is not None]
            if not params:
                block_eigenvalue.append(1.0)
                continue
            grads = [p.grad for p in params]
            v = self.normalize([torch.randn_like(p) for p in params])
            ev_cur, ev_prev, i = 1.0, 0.0, 0
            while i < self.max_iter and abs(ev_cur) > 0 and abs((ev_cur - ev_prev) / ev_cur) >= self.tol:
                ev_prev = ev_cur
                Hv = torch.autograd.grad(grads, params, grad_outputs=v, only_inputs=True, retain_graph=True)
                Hv = [self.nan_to_num(h.float()) for h in Hv]
                ev_cur = float(self.inner_product(Hv, v))
                v = self.normalize(Hv)
                v = [x / scale for x in v]
                i += 1
            block_eigenvalue.append(ev_cur * scale)
        mx = max(abs(e) for e in block_eigenvalue) or 1.0
        block_eigenvalue = [1.0 if e == 0 else abs(e) / mx for e in block_eigenvalue]
        ev_dict = {}
        for i, (block, ev) in enumerate(zip(layers, block_eigenvalue)):
            for p in block.parameters():
                ev_dict[id(p)] = (ev, i)
        return ev_dict
