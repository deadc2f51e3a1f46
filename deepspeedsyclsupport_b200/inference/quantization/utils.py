"""Group-wise int8 / int4 weight quantization for post-init weight-only quantized inference.

Storage format == the native quantizer's (csrc/quant/quantizer.cu): payload int8 [n * bits / 8] (int4: two values per
byte, low nibble first; asymmetric values are stored biased by -128 / -8), params fp32 [groups, 2] = (scale, offset).
Host tensors use the torch restatement below, device tensors the kernels (`ops/quantizer.py`).
(reference: deepspeed/inference/quantization/utils.py Quantizer / DeQuantizer)"""
import torch

from ...ops import quantizer as Q


def _to_last(t, dim):
    """group_dim -> last axis (groups must be contiguous runs for the kernels)"""
    return t if dim in (-1, t.dim() - 1) else t.transpose(dim, -1)


def quantize_host(x, groups, bits, symmetric):
    q, params = Q.int_quantize_ref(x, groups, bits, symmetric)
    q = q.to(torch.int32)
    if not symmetric:
        q = q - (128 if bits == 8 else 8)
    if bits == 8:
        payload = q.to(torch.int8).reshape(-1)
    else:
        q = q.reshape(-1, 2)
        payload = ((q[:, 0] & 0xF) | ((q[:, 1] & 0xF) << 4)).to(torch.uint8).view(torch.int8)
    return payload, params.float()


def dequantize_host(payload, params, bits, symmetric, dtype):
    groups = params.shape[0]
    if bits == 8:
        q = payload.to(torch.int32)
        if not symmetric:
            q = q + 128
    else:
        b = payload.view(torch.uint8).to(torch.int32)
        lo, hi = ((b & 0xF) ^ 8) - 8, (((b >> 4) & 0xF) ^ 8) - 8
        q = torch.stack([lo, hi], 1).reshape(-1)
        if not symmetric:
            q = q + 8
    q = q.view(groups, -1).float()
    return (q * params[:, 0:1] + params[:, 1:2]).to(dtype).reshape(-1)


class Quantizer(object):
    """config: {'num_bits': 4 | 8, 'group_size': int, 'group_dim': int, 'symmetric': bool}"""

    def __init__(self, config):
        self.config = config
        assert config["num_bits"] in (4, 8), "Only INT4 and INT8 quantization is supported."

    def quantize(self, tensor):
        """-> (payload int8 [numel * bits / 8], params fp32 [groups, 2], shape of the (group-dim-last) tensor)"""
        c = self.config
        t = _to_last(tensor.detach(), c.get("group_dim", 1)).contiguous()
        gs = c["group_size"]
        assert t.shape[-1] % gs == 0, f"Tensor shape: {tuple(tensor.shape)} quantization config {c}"
        groups = t.numel() // gs
        if t.is_cuda:
            payload, params = Q.int_quantize(t, groups, c["num_bits"], c.get("symmetric", False))
        else:
            payload, params = quantize_host(t, groups, c["num_bits"], c.get("symmetric", False))
        return payload, params, tuple(t.shape)


class DeQuantizer(object):

    def __init__(self, config, dtype):
        self.config, self.dtype = config, dtype

    def dequantize(self, payload, params, shape):
        c = self.config
        if payload.is_cuda:
            flat = Q.int_dequantize(payload, params, params.shape[0], c["num_bits"], c.get("symmetric", False), self.dtype)
        else:
            flat = dequantize_host(payload, params, c["num_bits"], c.get("symmetric", False), self.dtype)
        t = flat.view(shape)
        return _to_last(t, c.get("group_dim", 1)) if c.get("group_dim", 1) not in (-1, len(shape) - 1) else t


def recursive_setattr(model, module_name, module):
    parts = module_name.split(".")
    obj = model
    for p in parts[:-1]:
        obj = getattr(obj, p)
    setattr(obj, parts[-1], module)


# ---------------------------------------------------------------------------------------------------------------------
# fp16-on-host safe rounding helpers and the "one flat tensor per quantized weight" packing used when a quantized
# weight has to travel through code that expects a single tensor per parameter (state dicts, ZeRO-3 partitioning).
# Parity: inference/quantization/utils.py:24-40 (tensor_clamp / tensor_round), 176-288 (compat param + wrappers).
# ---------------------------------------------------------------------------------------------------------------------
import functools

import torch
from torch import nn


def _host_half(t):
    return t.device.type == "cpu" and t.dtype == torch.float16


def tensor_clamp(tensor, min, max):
    if _host_half(tensor):
        return tensor.float().clamp_(min, max).to(torch.float16)
    return tensor.clamp_(min, max)


def tensor_round(tensor):
    if _host_half(tensor):
        return tensor.float().round_().to(torch.float16)
    return tensor.round_()


def concat_to_compat_param(payload, params, shape, return_param=True):
    """Pack a quantized weight (``payload`` int8 bytes, ``params`` fp32 [groups, 2]) into one flat uint8 tensor.
    ``result.deconcat(result)`` gives back ``(payload, params, shape)`` as views."""
    n_payload, params_shape = payload.numel() * payload.element_size(), tuple(params.shape)
    flat = torch.cat([payload.contiguous().view(torch.uint8).flatten(),
                      params.contiguous().float().view(torch.uint8).flatten()])
    if return_param:
        flat = nn.Parameter(flat, requires_grad=False)
    payload_dtype, shape = payload.dtype, tuple(shape)

    def deconcat(t):
        data = t.data if isinstance(t, nn.Parameter) else t
        p = data.narrow(0, 0, n_payload).view(payload_dtype)
        s = data.narrow(0, n_payload, data.numel() - n_payload).view(torch.float32).view(params_shape)
        return p, s, shape

    flat.deconcat = deconcat
    return flat


def _quantize_param(param, quant_config):
    """Replace ``param.data`` by its packed quantized form and attach what is needed to undo it."""
    assert not getattr(param, "weight_quantized", False), "Parameter has already been quantized."
    quantizer, dequantizer = Quantizer(quant_config), DeQuantizer(quant_config, param.dtype)
    packed = concat_to_compat_param(*quantizer.quantize(param.data), return_param=False)
    param.deconcat = packed.deconcat
    param.requires_grad_(False)  # a packed integer weight is frozen
    param.data = packed
    param.quantizer, param.dequantizer = quantizer, dequantizer
    param.weight_quantized = True


def wrap_quantized_functional(f):
    """Make ``f(input, weight, ...)`` (``F.linear`` / ``F.embedding``) accept a packed quantized weight by
    de-quantizing it for the duration of the call."""

    @functools.wraps(f)
    def wrapper(input, weight, *args, **kwargs):
        if getattr(weight, "weight_quantized", False):
            payload, params, shape = weight.deconcat(weight)
            return f(input, weight.dequantizer.dequantize(payload, params, shape), *args, **kwargs)
        return f(input, weight, *args, **kwargs)

    return wrapper


def wrap_load_from_state_dict(f):
    """Make ``Module._load_from_state_dict`` quantize an incoming full-precision ``weight`` entry when the module's
    weight is already held packed; the caller's state dict is left as it was."""

    @functools.wraps(f)
    def wrapper(model, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        weight = getattr(model, "weight", None)
        key, saved = prefix + "weight", None
        if weight is not None and getattr(weight, "weight_quantized", False) and key in state_dict \
                and not getattr(weight, "state_dict_quantized", False) and state_dict[key].dtype != torch.uint8:
            weight.state_dict_quantized = True
            saved = state_dict[key]
            state_dict[key] = concat_to_compat_param(*weight.quantizer.quantize(saved), return_param=False)
        try:
            f(model, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs)
        finally:
            if saved is not None:
                state_dict[key] = saved
                del weight.state_dict_quantized

    return wrapper
