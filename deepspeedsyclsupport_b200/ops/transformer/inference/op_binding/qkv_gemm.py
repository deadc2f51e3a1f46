"""(parity: op_binding/qkv_gemm.py)  Input norm fused with the QKV projection: returns (qkv, normed input)."""
import torch

from ..ds_mlp import block_norm
from .base import BaseOp, linear_nt


class QKVGemmOp(BaseOp):

    def forward(self, input: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor):
        normed = block_norm(input, gamma, beta, self.config) if self.config.pre_layer_norm else input
        return linear_nt(normed, weight, bias), normed
