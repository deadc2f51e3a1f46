"""(parity: inference/v2/kernels/cutlass_ops/{moe_gemm,mixed_gemm}; the reference links external CUTLASS kernels.)
MoEGEMM is the repo's grouped tcgen05 GEMM (csrc/gemm/gemm_grouped.cu).  MixedGEMM / MixedMoEGEMM (16-bit activations x 4/8-bit
weights) dequantise group-wise into a 16-bit scratch and run the dense / grouped GEMM: weight-only quantisation saves HBM
capacity here, the W8A8 path (fp8 tensor cores, modules' fp8_linear) is the one that saves time."""
from typing import Optional

import torch

from .....ops import fused_ops as FO
from .....ops.gemm import gemm, moe_gemm
from ...inference_utils import ActivationType, DtypeEnum, activation_name, is_gated
from ..ds_kernel import DSKernelBase


def _act(out, x, bias, act_fn):
    if bias is not None:
        x = x + bias
    name = activation_name(act_fn)
    out.copy_(FO.gated_act(x, name) if is_gated(act_fn) else (x if ActivationType(act_fn) == ActivationType.IDENTITY else FO.bias_act(x, None, name)))
    return out


class MoEGEMM(DSKernelBase):
    supported_dtypes = (DtypeEnum.fp16, DtypeEnum.bf16)
    supported_act_fns = (ActivationType.GELU, ActivationType.SILU, ActivationType.RELU, ActivationType.IDENTITY)

    def __init__(self, fp_dtype, act_fn: ActivationType) -> None:
        if DtypeEnum(fp_dtype) not in self.supported_dtypes:
            raise ValueError(f"Unsupported data type: {fp_dtype}, supported_dtypes are {self.supported_dtypes}")
        if ActivationType(act_fn) not in self.supported_act_fns:
            raise ValueError(f"Unsupported activation function: {act_fn}, supported are {self.supported_act_fns}")
        self.act_fn = ActivationType(act_fn)

    def __call__(self, ordered_output: torch.Tensor, ordered_input: torch.Tensor, weights: torch.Tensor, cumsum_rows: torch.Tensor,
                 ordered_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        """ordered_input [rows sorted by expert, in]; weights [E, out, in]; cumsum_rows [E] int64 inclusive (device tensor)."""
        y = moe_gemm(ordered_input, weights, cumsum_rows)
        if ordered_bias is not None:
            rows = torch.arange(y.shape[0], device=y.device)
            y = y + ordered_bias[torch.searchsorted(cumsum_rows, rows, right=True).clamp_(max=weights.shape[0] - 1)]
        return _act(ordered_output, y, None, self.act_fn)


def _dequant(weights, scales, num_bits, dtype):
    """weights: int8 [..., out, in] (8-bit) or uint8 [..., out, in/2] (two 4-bit values per byte, low nibble first);
    scales [..., out, in / group] -> [..., out, in] in `dtype`."""
    if num_bits == 4:
        lo, hi = (weights & 0xF).to(torch.int8) - 8, (weights >> 4).to(torch.int8) - 8
        w = torch.stack([lo, hi], dim=-1).flatten(-2)
    else:
        w = weights
    group = w.shape[-1] // scales.shape[-1]
    return (w.to(torch.float32).unflatten(-1, (scales.shape[-1], group)) * scales.to(torch.float32).unsqueeze(-1)).flatten(-2).to(dtype)


class MixedGEMM(DSKernelBase):
    supported_dtypes = (DtypeEnum.fp16, DtypeEnum.bf16)

    def __init__(self, fp_dtype, act_fn: ActivationType, num_bits: int) -> None:
        if DtypeEnum(fp_dtype) not in self.supported_dtypes:
            raise ValueError(f"Unsupported data type: {fp_dtype}, supported_dtypes are {self.supported_dtypes}")
        if num_bits not in (4, 8):
            raise ValueError(f"Unsupported num_bits: {num_bits}, supported num_bits are 4 and 8")
        self.dtype, self.act_fn, self.num_bits = DtypeEnum(fp_dtype).value, ActivationType(act_fn), num_bits

    def __call__(self, output: torch.Tensor, hidden_states: torch.Tensor, weights: torch.Tensor, scales: torch.Tensor,
                 bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = _dequant(weights, scales, self.num_bits, hidden_states.dtype)
        x = hidden_states.reshape(-1, hidden_states.shape[-1])
        y = gemm(x, w) if (x.is_cuda and x.dtype == torch.bfloat16) else torch.nn.functional.linear(x, w)
        return _act(output, y.view(*hidden_states.shape[:-1], -1), bias, self.act_fn)


class MixedMoEGEMM(MixedGEMM):

    def __call__(self, ordered_output: torch.Tensor, ordered_input: torch.Tensor, weights: torch.Tensor, scales: torch.Tensor,
                 total_rows_before_expert: torch.Tensor, ordered_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = _dequant(weights, scales, self.num_bits, ordered_input.dtype)
        return MoEGEMM.__call__(self, ordered_output, ordered_input, w, total_rows_before_expert, ordered_bias)
