"""Local expert container (reference ``moe/experts.py:13``): ``num_local_experts`` deep copies of an expert
module; parameters are tagged ``allreduce=False`` + ``group_name`` so ZeRO/DP treat them as expert-parallel."""
import copy

import torch
from torch import nn


class Experts(nn.Module):

    def __init__(self, expert, num_local_experts=1, expert_group_name=None):
        super().__init__()
        self.deepspeed_experts = nn.ModuleList([copy.deepcopy(expert) for _ in range(num_local_experts)])
        self.num_local_experts = num_local_experts
        for e in self.deepspeed_experts:
            for p in e.parameters():
                p.allreduce = False
                p.group_name = expert_group_name

    def forward(self, inputs):
        # inputs: [E_local, tokens, M]
        outs = []
        for chunk, expert in zip(inputs.unbind(0), self.deepspeed_experts):
            out = expert(chunk)
            if isinstance(out, tuple):
                out = out[0]
            outs.append(out)
        return torch.stack(outs, dim=0)
