"""Muon (momentum + Newton-Schulz orthogonalised update for 2-D weights) with an auxiliary AdamW for everything else.

Reference: ``runtime/zero/muon/original_muon.py`` (``zeropower_via_newtonschulz5`` :37, Gram variant :69,
``muon_update`` :146, ``MuonWithAuxAdam`` :275) and ``muon_optimizer.py`` (under ZeRO the orthogonalisation is done by
the ZeRO optimizer on whole matrices and the inner optimizer only applies ``p -= lr * (update + wd * p)`` to its flat
shard). Same split here: :func:`muon_update` is called by ``ZeroShardedOptimizer._apply_muon`` on gathered matrices.
The Newton-Schulz chain is 15 small GEMMs per matrix, run in bf16/fp16 on the tensor cores."""
import torch

NS_METHODS = {"standard", "gram"}
_NS_COEFFS = (3.4445, -4.7750, 2.0315)


def _ns_dtype(t, prefer):
    return prefer if t.is_cuda else torch.float32


def zeropower_via_newtonschulz5(G, steps: int):
    """Quintic Newton-Schulz iteration pushing the singular values of ``G`` towards 1 (``U S' V^T``, S' in ~[0.5, 1.5])."""
    assert G.ndim >= 2
    a, b, c = _NS_COEFFS
    X = G.to(_ns_dtype(G, torch.bfloat16))
    tall = G.size(-2) > G.size(-1)
    if tall:
        X = X.mT
    X = X / (X.norm(dim=(-2, -1), keepdim=True) + 1e-7)
    for _ in range(steps):
        A = X @ X.mT
        B = torch.addmm(A, A, A, beta=b, alpha=c) if A.ndim == 2 else b * A + c * (A @ A)
        X = torch.addmm(X, B, X, beta=a) if X.ndim == 2 else a * X + B @ X
    return X.mT if tall else X


def zeropower_via_gram_newtonschulz(G, steps: int):
    """Same polynomial iterated on the small Gram matrix ``R = X X^T`` (n x n) instead of the n x m iterate: far fewer
    FLOPs for wide matrices. fp16 with a restart at iteration 2 for stability; square inputs use the standard form."""
    assert G.ndim >= 2
    a, b, c = _NS_COEFFS
    X = G.to(_ns_dtype(G, torch.float16))
    tall = G.size(-2) > G.size(-1)
    if tall:
        X = X.mT
    n, m = X.size(-2), X.size(-1)
    X = X / (X.norm(dim=(-2, -1), keepdim=True) + 1e-7)
    if m <= n:
        for _ in range(steps):
            A = X @ X.mT
            X = a * X + (b * A + c * A @ A) @ X
        return X.mT if tall else X
    R = X @ X.mT
    Q = None
    restart_at = 2
    for i in range(steps):
        if i == restart_at and i != 0:
            X = Q @ X
            R = X @ X.mT
            Q = None
        Z = b * R + c * R @ R
        if Q is None:
            Q = Z.clone()
            Q.diagonal(dim1=-2, dim2=-1).add_(a)
        else:
            Q = a * Q + Z @ Q
        if i < steps - 1 and (i + 1) != restart_at:
            RZ = a * R + Z @ R
            R = a * RZ + Z @ RZ
    return (X.mT @ Q.mT) if tall else (Q @ X)


def muon_update(grad, momentum, beta=0.95, ns_steps=5, nesterov=True, ns_method="gram", is_expert_group=False):
    """In-place momentum update, then the orthogonalised (and aspect-ratio scaled) step direction."""
    orig_dtype = grad.dtype
    momentum.lerp_(grad, 1 - beta)
    update = grad.lerp_(momentum, beta) if nesterov else momentum
    ns = zeropower_via_gram_newtonschulz if ns_method == "gram" else zeropower_via_newtonschulz5
    if is_expert_group:  # [experts, out, in]: batched orthogonalisation
        scale = max(1, update.size(-2) / update.size(-1))**0.5
        update = ns(update, steps=ns_steps) * scale
    else:
        if update.ndim == 4:  # conv filters
            update = update.view(len(update), -1)
        rows, cols = update.size(-2), update.size(-1)
        update = ns(update, steps=ns_steps) * max(1, rows / cols)**0.5
    return update.to(orig_dtype).view_as(grad) if update.numel() == grad.numel() else update.to(orig_dtype)


def adam_update(grad, buf1, buf2, step, betas, eps):
    buf1.lerp_(grad, 1 - betas[0])
    buf2.lerp_(grad.square(), 1 - betas[1])
    return (buf1 / (1 - betas[0]**step)) / ((buf2 / (1 - betas[1]**step)).sqrt() + eps)


def build_muon_param_groups(params, optimizer_parameters):
    """Split by ``param.use_muon`` (default: matrices yes, vectors/scalars no) into a muon and an adam group, mapping
    ``muon_lr`` / ``adam_lr`` onto each group's ``lr`` (reference engine.py:2200-2233)."""
    flat = []
    for item in params:
        flat.extend(item["params"] if isinstance(item, dict) else [item])
    flat = [p for p in flat if p.requires_grad]
    op = dict(optimizer_parameters or {})
    muon = [p for p in flat if getattr(p, "use_muon", p.ndim >= 2)]
    rest = [p for p in flat if not getattr(p, "use_muon", p.ndim >= 2)]
    groups = []
    if muon:
        g = {k: op[k] for k in ("lr", "momentum", "weight_decay", "ns_method") if k in op}
        if "muon_lr" in op:
            g["lr"] = op["muon_lr"]
        groups.append(dict(params=muon, use_muon=True, name="muon-params", **g))
    if rest:
        g = {k: op[k] for k in ("lr", "betas", "eps", "weight_decay") if k in op}
        if "adam_lr" in op:
            g["lr"] = op["adam_lr"]
        groups.append(dict(params=rest, use_muon=False, name="adam-params", **g))
    return groups


class MuonWithAuxAdam(torch.optim.Optimizer):
    """``param_groups`` carry ``use_muon``. ``zero_mode`` (set by the ZeRO wrapper): parameters are flat shards whose
    ``.grad`` already holds the orthogonalised update, so the muon groups only apply it."""

    def __init__(self, param_groups, zero_mode=False):
        for g in param_groups:
            assert "use_muon" in g
            if g["use_muon"]:
                g.setdefault("lr", 0.02), g.setdefault("momentum", 0.95), g.setdefault("weight_decay", 0)
                g.setdefault("ns_method", "gram")
                assert g["ns_method"] in NS_METHODS
            else:
                g.setdefault("lr", 3e-4), g.setdefault("betas", (0.9, 0.95)), g.setdefault("eps", 1e-10)
                g.setdefault("weight_decay", 0)
        super().__init__(param_groups, dict())
        self.zero_mode = zero_mode
        self.needs_full_matrix_update = True  # tells ZeroShardedOptimizer to run muon_update on whole matrices

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if group["use_muon"]:
                    if self.zero_mode:
                        update = p.grad.reshape(p.shape)
                    else:
                        st = self.state[p]
                        if "momentum_buffer" not in st:
                            st["momentum_buffer"] = torch.zeros_like(p)
                        update = muon_update(p.grad, st["momentum_buffer"], beta=group["momentum"],
                                             ns_method=group["ns_method"]).reshape(p.shape)
                    p.mul_(1 - group["lr"] * group["weight_decay"])
                    p.add_(update.to(p.dtype), alpha=-group["lr"])
                else:
                    st = self.state[p]
                    if len(st) == 0:
                        st["exp_avg"], st["exp_avg_sq"], st["step"] = torch.zeros_like(p), torch.zeros_like(p), 0
                    st["step"] += 1
                    update = adam_update(p.grad, st["exp_avg"], st["exp_avg_sq"], st["step"], group["betas"],
                                         group["eps"])
                    p.mul_(1 - group["lr"] * group["weight_decay"])
                    p.add_(update, alpha=-group["lr"])
        return loss


class Muon(MuonWithAuxAdam):
    """``Muon(params, lr=..., momentum=..., muon_lr=..., adam_lr=..., betas=..., eps=..., weight_decay=...)``: the
    ds_config form (``optimizer.type = "Muon"``)."""

    def __init__(self, params, **optimizer_parameters):
        params = list(params)
        if params and isinstance(params[0], dict) and "use_muon" in params[0]:
            groups = params
        else:
            groups = build_muon_param_groups(params, optimizer_parameters)
        super().__init__(groups)


SingleDeviceMuon = Muon
SingleDeviceMuonWithAuxAdam = MuonWithAuxAdam
