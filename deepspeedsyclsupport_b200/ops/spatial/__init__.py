"""Spatial (diffusers) ops: channels-last bias-add family (parity: csrc/spatial + op_builder/spatial_inference.py:
nhwc_bias_add, nhwc_bias_add_add, nhwc_bias_add_bias_add)."""
import torch

from .. import native

native.register_sigs({"dsb200_nhwc_bias_add": "ppppp l ii p"})
_DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


def nhwc_bias_add(activation, bias, other=None, other_bias=None):
    """activation (+ other) are channels-last tensors whose innermost *memory* dimension is the channel."""
    channels = bias.numel()
    usable = activation.is_cuda and activation.dtype in _DT and (channels * activation.element_size()) % 16 == 0
    cl = activation.dim() == 4 and activation.is_contiguous(memory_format=torch.channels_last)
    if not usable or not (cl or (activation.is_contiguous() and activation.shape[-1] == channels)):
        shape = (1, -1, 1, 1) if activation.dim() == 4 and activation.shape[1] == channels else (-1, )
        out = activation + bias.view(shape)
        if other is not None:
            out = out + other
        if other_bias is not None:
            out = out + other_bias.view(shape)
        return out
    if other is not None:
        other = other.contiguous(memory_format=torch.channels_last) if cl else other.contiguous()
    out = torch.empty_like(activation)
    native.check(native.lib().dsb200_nhwc_bias_add(native.ptr(activation), native.ptr(bias.contiguous()), native.ptr(other),
                                                   native.ptr(other_bias.contiguous() if other_bias is not None else None), native.ptr(out),
                                                   activation.numel(), channels, _DT[activation.dtype], native.stream_ptr()))
    return out


def nhwc_bias_add_add(activation, bias, other):
    return nhwc_bias_add(activation, bias, other=other)


def nhwc_bias_add_bias_add(activation, bias, other, other_bias):
    return nhwc_bias_add(activation, bias, other=other, other_bias=other_bias)
