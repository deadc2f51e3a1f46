"""(parity: op_binding/mlp_gemm.py)  residual add + post-attention norm + up projection + activation + down projection.
Returns (mlp output, residual after the add)."""
from typing import Optional

import torch

from .....utils.types import GATED_ACTIVATION_TYPES
from .... import fused_ops as FO
from ..ds_mlp import _ACT, block_norm
from .base import BaseOp, linear_nt


class MLPGemmOp(BaseOp):

    def forward(self, input: torch.Tensor, residual: torch.Tensor, weight_interm: torch.Tensor, weight_out: torch.Tensor,
                input_bias: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None, gamma: Optional[torch.Tensor] = None,
                beta: Optional[torch.Tensor] = None):
        cfg = self.config
        residual_add = input + residual + (input_bias if input_bias is not None else 0)
        h = block_norm(residual_add, gamma, beta, cfg)
        u = linear_nt(h, weight_interm)
        act = _ACT[cfg.mlp_act_func_type]
        if cfg.mlp_act_func_type in GATED_ACTIVATION_TYPES:
            a = FO.gated_act(u if bias is None else u + bias, act)
        else:
            a = FO.bias_act(u, bias, act)
        return linear_nt(a, weight_out), residual_add
