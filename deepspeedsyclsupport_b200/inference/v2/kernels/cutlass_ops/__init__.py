"""(parity: inference/v2/kernels/cutlass_ops/{moe_gemm,mixed_gemm}; the reference links external CUTLASS kernels.)
MoEGEMM is the repo's grouped tcgen05 GEMM (csrc/gemm/gemm_grouped.cu).  MixedGEMM (16-bit activations x 4/8-bit weights):
decode-sized calls (<= 128 rows) run the mixed-input weight-streaming kernel (csrc/gemm/gemm_skinny.cu: raw integer tiles by TMA,
dequantised to bf16 in shared memory by the epilogue warps, tcgen05 MMA) -- HBM traffic is the quantised bytes; prefill-sized
calls and MixedMoEGEMM dequantise group-wise into a 16-bit scratch and run the dense / grouped GEMM (compute-bound there)."""
from typing import Optional

import torch

from .....ops import fused_ops as FO
from .....ops import native
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


native.register_sigs({"dsb200_gemm_skinny_mixed": "ppppp iii ii iii p"})


def _mixed_skinny(x, weights, scales, num_bits):
    """x [M <= 128, K] bf16 x quantised weights [N, K] -> [M, N] bf16 on the weight-streaming kernel (csrc/gemm/gemm_skinny.cu,
    skinny_mixed_kernel), or None when the shapes / dtypes are outside it."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and weights.dim() == 2 and x.shape[0] <= 128 and x.stride(1) == 1):
        return None
    M, K = x.shape
    N = weights.shape[0]
    n_groups = scales.shape[-1]
    if K % 64 or K % n_groups or (K // n_groups) % 64 or N % 4 or x.stride(0) % 8 or weights.shape[1] * (2 if num_bits == 4 else 1) != K:
        return None
    sc = scales if (scales.dtype == torch.float32 and scales.is_contiguous()) else scales.float().contiguous()
    wq = weights if weights.is_contiguous() else weights.contiguous()
    tiles, kb = -(-N // 128), K // 64
    splits = max(1, min(148 // tiles, kb // (32 if tiles >= 32 else 8))) if tiles < 148 else 1  # as ops.gemm.gemm
    out = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    ws = torch.empty(splits * M * N, dtype=torch.float32, device=x.device) if splits > 1 else None
    rc = native.lib().dsb200_gemm_skinny_mixed(native.ptr(x), native.ptr(wq), native.ptr(sc), native.ptr(out), native.ptr(ws), M, N, K,
                                               x.stride(0), N, int(num_bits), K // n_groups, splits, native.stream_ptr())
    native.check(rc)
    return out


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
        x = hidden_states.reshape(-1, hidden_states.shape[-1])
        y = _mixed_skinny(x, weights, scales, self.num_bits)
        if y is not None:  # decode-sized: the quantised bytes are streamed once and dequantised in shared memory
            return _act(output, y.view(*hidden_states.shape[:-1], -1), bias, self.act_fn)
        # prefill-sized (compute-bound): group-wise dequantisation into a 16-bit scratch, then the dense tcgen05 GEMM
        w = _dequant(weights, scales, self.num_bits, hidden_states.dtype)
        y = gemm(x, w) if (x.is_cuda and x.dtype == torch.bfloat16) else torch.nn.functional.linear(x, w)
        return _act(output, y.view(*hidden_states.shape[:-1], -1), bias, self.act_fn)


class MixedMoEGEMM(MixedGEMM):

    def __call__(self, ordered_output: torch.Tensor, ordered_input: torch.Tensor, weights: torch.Tensor, scales: torch.Tensor,
                 total_rows_before_expert: torch.Tensor, ordered_bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        w = _dequant(weights, scales, self.num_bits, ordered_input.dtype)
        return MoEGEMM.__call__(self, ordered_output, ordered_input, w, total_rows_before_expert, ordered_bias)
