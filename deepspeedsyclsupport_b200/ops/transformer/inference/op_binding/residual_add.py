"""(parity: op_binding/residual_add.py)  The block's closing sum.  With tensor parallelism the biases and the residual are
added on every rank *before* the all-reduce of the partial outputs, hence the 1 / mp_size factors."""
from typing import Optional

import torch

from .base import BaseOp


class ResidualAddOp(BaseOp):

    def forward(self, hidden_state: torch.Tensor, residual: torch.Tensor, add_bias: bool, attention_output: Optional[torch.Tensor] = None,
                residual_add: Optional[torch.Tensor] = None, attention_bias: Optional[torch.Tensor] = None,
                final_bias: Optional[torch.Tensor] = None):
        cfg = self.config
        inv = 1.0 / cfg.mp_size
        if not cfg.mlp_after_attn:
            # parallel block: hidden (MLP) + attention + residual, biases once per rank-sum
            out = hidden_state + (attention_output if attention_output is not None else 0) + residual * inv
            if add_bias and final_bias is not None:
                out = out + final_bias * inv
            if attention_bias is not None:
                out = out + attention_bias * inv
            return out
        base = residual_add if (not cfg.pre_layer_norm and residual_add is not None) else residual
        out = hidden_state + base * inv
        if add_bias and final_bias is not None:
            out = out + final_bias * inv
        return out
