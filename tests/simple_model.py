"""Tiny models/data for engine tests (reference ``tests/unit/simple_model.py``)."""
import torch
import torch.nn as nn


class SimpleModel(nn.Module):

    def __init__(self, hidden_dim=16, nlayers=2, empty_grad=False):
        super().__init__()
        self.linears = nn.ModuleList([nn.Linear(hidden_dim, hidden_dim) for _ in range(nlayers)])
        self.norm = nn.LayerNorm(hidden_dim)
        if empty_grad:
            self.unused = nn.Linear(hidden_dim, hidden_dim)
        self.loss = nn.CrossEntropyLoss()

    def forward(self, x, y):
        for l in self.linears:
            x = torch.relu(l(x)) + x
        x = self.norm(x)
        return self.loss(x, y)


def random_batches(n, batch, hidden_dim, device, dtype=torch.float32, seed=0):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        x = torch.randn(batch, hidden_dim, generator=g).to(device=device, dtype=dtype)
        y = torch.randint(0, hidden_dim, (batch, ), generator=g).to(device)
        out.append((x, y))
    return out
