"""(parity: op_binding/gelu_gemm.py)  act(x W1^T + b) W2^T for blocks whose MLP reads the already normalised input (GPT-J)."""
import torch

from .... import fused_ops as FO
from ..ds_mlp import _ACT
from .base import BaseOp, linear_nt


class GELUGemmOp(BaseOp):

    def forward(self, input: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, weight_out: torch.Tensor):
        a = FO.bias_act(linear_nt(input, weight), bias, _ACT[self.config.mlp_act_func_type])
        return linear_nt(a, weight_out)
