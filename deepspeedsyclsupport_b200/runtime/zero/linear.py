"""Memory-lean linear for ZeRO stage 3 (parity: deepspeed/runtime/zero/linear.py:43-178).

torch's linear keeps `weight.t()` views alive in the autograd graph, which pins the gathered full weight until backward; this
Function saves the weight *parameter* itself (a view into the flat unit buffer that is re-gathered before the backward of its
module), so the gathered storage can be released after forward.  bf16 CUDA operands run on the tcgen05 GEMM."""
import math

import torch
from torch import Tensor
from torch.nn import init
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter


def _mm(a, b_t):
    """a [M, K] @ b_t[N, K]^T"""
    if a.is_cuda and a.dtype == torch.bfloat16 and b_t.dtype == torch.bfloat16 and a.shape[-1] % 8 == 0 and a.is_contiguous() and b_t.is_contiguous():
        from ...ops.gemm import gemm
        return gemm(a, b_t)
    return a @ b_t.t()


class LinearFunctionForZeroStage3(torch.autograd.Function):

    @staticmethod
    def forward(ctx, input, weight, bias=None):
        ctx.save_for_backward(input, weight, bias)
        x2 = input.reshape(-1, input.shape[-1])
        out = _mm(x2, weight).view(*input.shape[:-1], weight.shape[0])
        if bias is not None:
            out = out + bias
        return out

    @staticmethod
    def backward(ctx, grad_output):
        input, weight, bias = ctx.saved_tensors
        grad_input = grad_weight = grad_bias = None
        go = grad_output.reshape(-1, grad_output.shape[-1])
        if ctx.needs_input_grad[0]:
            grad_input = (go @ weight).view(input.shape)
        if ctx.needs_input_grad[1]:
            grad_weight = go.t() @ input.reshape(-1, input.shape[-1])
        if bias is not None and ctx.needs_input_grad[2]:
            grad_bias = go.sum(0)
        return grad_input, grad_weight, grad_bias


def zero3_linear_wrap(input, weight, bias=None):
    return LinearFunctionForZeroStage3.apply(input, weight, bias)


class LinearModuleForZeroStage3(Module):
    """nn.Linear with the memory-lean function (installed for nn.Linear under `zero.Init(mem_efficient_linear=True)`)."""
    __constants__ = ["in_features", "out_features"]
    in_features: int
    out_features: int
    weight: Tensor

    def __init__(self, in_features: int, out_features: int, bias: bool = True) -> None:
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = Parameter(torch.empty(out_features, in_features))
        if bias:
            self.bias = Parameter(torch.empty(out_features))
        else:
            self.register_parameter("bias", None)
        self.reset_parameters()

    def reset_parameters(self) -> None:
        init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            fan_in, _ = init._calculate_fan_in_and_fan_out(self.weight)
            bound = 1 / math.sqrt(fan_in)
            init.uniform_(self.bias, -bound, bound)

    def forward(self, input: Tensor) -> Tensor:
        return LinearFunctionForZeroStage3.apply(input, self.weight, self.bias)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}"
