"""(parity: deepspeed/ops/transformer/inference/bias_add.py)  Channels-last bias add (+ optional second activation and bias) of
the diffusers blocks; the kernel is csrc/spatial/bias_add.cu."""
from typing import Optional

import torch

from ...spatial import nhwc_bias_add as _nhwc_bias_add


def nhwc_bias_add(activation: torch.Tensor, bias: torch.Tensor, other: Optional[torch.Tensor] = None,
                  other_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    return _nhwc_bias_add(activation, bias, other=other, other_bias=other_bias)
