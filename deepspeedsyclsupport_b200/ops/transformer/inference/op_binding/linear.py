"""(parity: op_binding/linear.py)  Plain projection used by the diffusers attention (fused QKV with optional bias)."""
import torch

from .base import BaseOp, linear_nt


class LinearOp(BaseOp):

    def forward(self, input: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor = None, add_bias: bool = False, do_flash_attn: bool = False,
                num_heads: int = 1, external_cache: bool = None, num_layers: int = None):
        w = weight if not self.config.transposed_mode else weight.t()
        return linear_nt(input, w, bias if add_bias else None)
