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
