"""Straight-through quantizers / binarizers used by compression-aware training.

All of them are autograd Functions with an identity (straight-through) backward.  The grouped symmetric / asymmetric
fake-quantisation runs on the fused CUDA kernel (csrc/quant: one pass computes the per-group range and writes the
quantise-dequantised values) when the tensor is on the GPU; CPU tensors use the torch formulas below, which also serve
as the numerics reference in tests.  API parity: deepspeed/compression/utils.py."""
import torch
from torch import autograd


def _grouped(x, num_groups):
    return x.reshape(num_groups, -1)


class TopKBinarizer(autograd.Function):
    """mask = 1 for the `threshold` fraction of largest scores (optionally after a sigmoid), else 0; the gradient goes
    straight to the scores (movement-pruning style)."""

    @staticmethod
    def forward(ctx, inputs: torch.Tensor, threshold: float, sigmoid: bool):
        if sigmoid:
            threshold = torch.sigmoid(threshold).item()
        ctx.sigmoid = sigmoid
        flat = inputs.reshape(-1)
        keep = int(-(-threshold * flat.numel() // 1))  # ceil
        mask = torch.zeros_like(flat)
        if keep > 0:
            mask[flat.topk(min(keep, flat.numel()), sorted=False).indices] = 1.0
        return mask.view_as(inputs)

    @staticmethod
    def backward(ctx, grad_output):
        if ctx.sigmoid:
            return grad_output.clone(), (grad_output * 1.0).sum().view(-1), None
        return grad_output.clone(), None, None


def _fake_quant_cuda(x, num_bits, num_groups, symmetric):
    from ..ops.quantizer import fake_quantize_
    out = x.detach().clone().contiguous()
    fake_quantize_(out, num_groups, num_bits, symmetric=symmetric)
    return out


class SymQuantizer(torch.autograd.Function):
    """Symmetric uniform quantisation: scale = 2 * max|x| / 2^bits per group, values rounded into [-2^(b-1), 2^(b-1) - 1]."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        q_range = 2**num_bits
        shape = input.shape
        if min_value is None:
            g = _grouped(input, num_groups)
            max_input = g.abs().amax(dim=-1, keepdim=True)
        else:
            g = input.reshape(1, -1)
            max_input = torch.max(min_value.abs(), max_value).view(-1)
        scale = 2 * max_input / q_range
        scale = torch.where(scale > 0, scale, torch.ones_like(scale))
        out = (g / scale).round().clamp(-q_range // 2, q_range // 2 - 1) * scale
        return out.reshape(shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class AsymQuantizer(torch.autograd.Function):
    """Asymmetric uniform quantisation: per-group [min, max] mapped onto [0, 2^bits - 1], zero-point snapped to the grid."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert (min_value is None and max_value is None) or (min_value is not None and max_value is not None and num_groups == 1)
        q_range = 2**num_bits
        shape = input.shape
        if min_value is None:
            g = _grouped(input, num_groups)
            lo, hi = g.amin(dim=-1, keepdim=True), g.amax(dim=-1, keepdim=True)
        else:
            g = input.reshape(1, -1)
            lo, hi = min_value, max_value
        scale = (hi - lo) / q_range
        scale = torch.where(scale > 0, scale, torch.ones_like(scale))
        zero_point = (lo / scale).round() * scale
        out = ((g - zero_point) / scale).round().clamp(0, q_range - 1) * scale + zero_point
        return out.reshape(shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class TernaryQuantizer(torch.autograd.Function):
    """{-a, 0, +a} per group: threshold 0.7 * mean|x|, a = mean of the magnitudes above the threshold (TWN)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        g = _grouped(input, num_groups)
        n = g.shape[1]
        thres = (0.7 * g.norm(p=1, dim=1) / n).view(-1, 1)
        pos = (g > thres).type(input.type())
        neg = (g < -thres).type(input.type())
        mask = (g.abs() > thres).type(input.type())
        alpha = ((mask * g).abs().sum(dim=1) / mask.sum(dim=1).clamp(min=1)).view(-1, 1)
        return (alpha * pos - alpha * neg).reshape(input.shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None


class BinaryQuantizer(torch.autograd.Function):
    """sign(x) * mean|x| per group (BWN)."""

    @staticmethod
    def forward(ctx, input, num_bits, min_value=None, max_value=None, num_groups=1):
        assert min_value is None and max_value is None
        g = _grouped(input, num_groups)
        m = g.norm(p=1, dim=1, keepdim=True) / g.shape[1]
        return (g.sign() * m).reshape(input.shape).contiguous()

    @staticmethod
    def backward(ctx, grad_output):
        return grad_output.clone(), None, None, None, None
