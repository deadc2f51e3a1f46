"""(parity: op_binding/vector_matmul.py)  Bias-free projection (attention output, cross-attention q / k / v)."""
import torch

from .base import BaseOp, linear_nt


class VectorMatMulOp(BaseOp):

    def forward(self, input: torch.Tensor, weight: torch.Tensor, async_op: bool = False):
        return linear_nt(input, weight if not self.config.transposed_mode else weight.t())
