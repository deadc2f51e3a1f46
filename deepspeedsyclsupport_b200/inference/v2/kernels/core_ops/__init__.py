"""Dense building-block kernels as configured objects (parity: inference/v2/kernels/core_ops/{bias_activations, blas_kernels,
cuda_layer_norm, cuda_rms_norm, gated_activations}).  Each writes into the caller's output buffer, as the reference's do; the
device code is csrc/norm, csrc/fused (bias/gated activations) and the tcgen05 GEMM."""
from typing import Optional

import torch

from .....ops import fused_ops as FO
from .....ops.gemm import gemm
from ...inference_utils import ActivationType, DtypeEnum, activation_name, is_gated
from ..ds_kernel import DSKernelBase

_FP = (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32)


def _check_dtype(dtype, who):
    if DtypeEnum(dtype) not in _FP:
        raise ValueError(f"Unsupported data type: {dtype} for {who}, supported_dtypes are {[d.value for d in _FP]}")
    return DtypeEnum(dtype).value


def _check_channels(channels, dtype, who):
    """rows are moved as 16-byte vectors"""
    per_vec = 16 // torch.empty((), dtype=dtype).element_size()
    if channels % per_vec != 0:
        raise ValueError(f"channels must be divisible by {per_vec} for {who} ({dtype}), got {channels}")


class CUDABiasActivation(DSKernelBase):
    """In place: activation = act(activation + bias)."""

    def __init__(self, channels: int, dtype, act_fn: ActivationType) -> None:
        self.dtype = _check_dtype(dtype, "CUDABiasActivation")
        _check_channels(channels, self.dtype, "CUDABiasActivation")
        act_fn = ActivationType(act_fn)
        if act_fn not in (ActivationType.IDENTITY, ActivationType.GELU, ActivationType.RELU, ActivationType.SILU):
            raise ValueError(f"Unsupported activation function: {act_fn}")
        self.act = activation_name(act_fn)

    def __call__(self, activation: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        activation.copy_(FO.bias_act(activation, bias, self.act))
        return activation


class CUDAGatedActivation(DSKernelBase):
    """output[T, C] = act(input[T, :C] + bias[:C]) * (input[T, C:] + bias[C:])  ([gate | up] halves)."""

    def __init__(self, channels: int, fp_dtype, act_fn: ActivationType) -> None:
        self.dtype = _check_dtype(fp_dtype, "CUDAGatedActivation")
        _check_channels(channels, self.dtype, "CUDAGatedActivation")
        if not is_gated(act_fn):
            raise ValueError(f"Unsupported activation function: {act_fn}, supported are GEGLU, ReGLU, SiGLU")
        self.channels, self.act = channels, activation_name(act_fn)

    def __call__(self, output: torch.Tensor, input: torch.Tensor, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        assert input.shape[-1] == 2 * self.channels and output.shape[-1] == self.channels
        output.copy_(FO.gated_act(input if bias is None else input + bias, self.act))
        return output


class BlasLibLinear(DSKernelBase):
    """output = hidden_states @ weights^T (weights [out, in], as nn.Linear).  bf16 runs on the repo's tcgen05 GEMM."""

    def __init__(self, fp_dtype):
        self.dtype = _check_dtype(fp_dtype, "BlasLibLinear")

    def __call__(self, output: torch.Tensor, hidden_states: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
        x = hidden_states.reshape(-1, hidden_states.shape[-1])
        if x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0:
            gemm(x, weights, out=output.view(-1, output.shape[-1]))
        else:
            output.copy_(torch.nn.functional.linear(hidden_states, weights))
        return output


class _NormBase(DSKernelBase):

    def __init__(self, channels: int, fp_dtype, epsilon: float = 1e-5):
        self.dtype = _check_dtype(fp_dtype, type(self).__name__)
        _check_channels(channels, self.dtype, type(self).__name__)
        self.channels, self.eps = channels, epsilon


class CUDARMSNorm(_NormBase):

    def __call__(self, output_z: torch.Tensor, input_x: torch.Tensor, gamma: torch.Tensor) -> torch.Tensor:
        output_z.copy_(FO.rms_norm(input_x, gamma, self.eps))
        return output_z


class CUDARMSPreNorm(_NormBase):
    """residual_out = residual + hidden;  z_out = RMSNorm(residual_out) * gamma -- one pass."""

    def __call__(self, z_res: torch.Tensor, z_hid: torch.Tensor, x_res: torch.Tensor, x_hid: torch.Tensor, gamma: torch.Tensor):
        y, new_res = FO.rms_norm(x_hid, gamma, self.eps, x_res)
        z_res.copy_(new_res)
        z_hid.copy_(y)
        return z_res, z_hid


class CUDAFPLN(_NormBase):

    def __call__(self, output_z: torch.Tensor, input_x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor) -> torch.Tensor:
        output_z.copy_(FO.layer_norm(input_x, gamma, beta, self.eps))
        return output_z


class CUDAFPPreLN(_NormBase):
    """residual_out = residual + hidden;  z_out = LayerNorm(residual_out)."""

    def __call__(self, z_res: torch.Tensor, z_hid: torch.Tensor, x_res: torch.Tensor, x_hid: torch.Tensor, gamma: torch.Tensor,
                 beta: torch.Tensor):
        y, new_res = FO.layer_norm(x_hid, gamma, beta, self.eps, x_res)
        z_res.copy_(new_res)
        z_hid.copy_(y)
        return z_res, z_hid


class CUDAFPPostLN(_NormBase):
    """z = LayerNorm(residual + hidden)."""

    def __call__(self, output_z: torch.Tensor, input_x: torch.Tensor, input_residual: torch.Tensor, gamma: torch.Tensor,
                 beta: torch.Tensor) -> torch.Tensor:
        y, _ = FO.layer_norm(input_x, gamma, beta, self.eps, input_residual)
        output_z.copy_(y)
        return output_z
