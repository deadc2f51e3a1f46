"""Local experts container (parity: deepspeed/moe/experts.py:10-35)."""
import copy

import torch


class Experts(torch.nn.Module):

    def __init__(self, expert, num_local_experts=1, expert_group_name=None):
        super().__init__()
        self.deepspeed_experts = torch.nn.ModuleList([copy.deepcopy(expert) for _ in range(num_local_experts)])
        self.num_local_experts = num_local_experts
        # expert parameters are NOT all-reduced over the dense data-parallel group: they belong to the expert-dp group
        for expert in self.deepspeed_experts:
            for _, param in expert.named_parameters():
                param.allreduce = False
                param.group_name = expert_group_name

    def forward(self, inputs):
        """inputs: [ep, local_E, C, M] -> same shape; expert e consumes chunk e along dim 1."""
        chunks = inputs.chunk(self.num_local_experts, dim=1)
        expert_outputs = []
        for chunk, expert in zip(chunks, self.deepspeed_experts):
            out = expert(chunk)
            if isinstance(out, tuple):
                out = out[0]  # ignore bias terms of Megatron-style experts
            expert_outputs.append(out)
        return torch.cat(expert_outputs, dim=1)
