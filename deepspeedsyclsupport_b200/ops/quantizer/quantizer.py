"""Quantizer ops (csrc/quant/quantizer.cu) with PyTorch oracles.

fp8 block-scaled (e4m3 + fp32 scale / 128 elements): `quantize_fp8_block`, `dequantize_fp8_block`,
`dequant_reduce_fp8_block` -- the ZeRO++ wire format here.  int8/int4 group quantization and fake quantization keep
the reference's `Quantizer` / `ds_quantizer` API (deepspeed/ops/quantizer/quantizer.py, runtime/quantize.py)."""
import torch

from .. import native

native.register_sigs({
    "dsb200_fp8_block_quantize": "ppp l i p",
    "dsb200_fp8_block_dequantize": "ppp l i p",
    "dsb200_fp8_block_dequant_reduce": "ppp l i f i p",
    "dsb200_int_quantize": "ppp l iiii p",
    "dsb200_int_dequantize": "ppp l iiii p",
    "dsb200_fake_quantize": "p l iiii l i p",
})

FP8_BLOCK = 128
FP8_MAX = 448.0
_DT = {torch.bfloat16: 0, torch.float16: 1, torch.float32: 2}


def quantize_fp8_block_ref(x):
    xf = x.float().view(-1, FP8_BLOCK)
    amax = xf.abs().amax(dim=1)
    scale = torch.where(amax > 0, amax / FP8_MAX, torch.ones_like(amax))
    q = (xf / scale[:, None]).to(torch.float8_e4m3fn)
    return q.view(torch.uint8).view(-1), scale


def dequantize_fp8_block_ref(q, scales, dtype=torch.bfloat16):
    return (q.view(torch.float8_e4m3fn).float().view(-1, FP8_BLOCK) * scales[:, None]).to(dtype).view(-1)


def quantize_fp8_block(x, out=None):
    """x: flat tensor with numel % 128 == 0 -> (uint8 payload [n], fp32 scales [n/128]).  `out=(payload, scales)` writes
    into existing buffers (e.g. regions of a symmetric buffer that peers read)."""
    x = x.contiguous().view(-1)
    n = x.numel()
    assert n % FP8_BLOCK == 0, "pad to a multiple of 128 elements"
    if not x.is_cuda:
        q, s = quantize_fp8_block_ref(x)
        if out is not None:
            out[0].copy_(q)
            out[1].copy_(s)
            return out
        return q, s
    if out is not None:
        q, s = out
        assert q.numel() == n and s.numel() == n // FP8_BLOCK and q.dtype == torch.uint8 and s.dtype == torch.float32
    else:
        q = torch.empty(n, dtype=torch.uint8, device=x.device)
        s = torch.empty(n // FP8_BLOCK, dtype=torch.float32, device=x.device)
    rc = native.lib().dsb200_fp8_block_quantize(native.ptr(x), native.ptr(q), native.ptr(s), n, _DT[x.dtype], native.stream_ptr())
    native.check(rc)
    return q, s


def dequantize_fp8_block(q, scales, out=None, dtype=torch.bfloat16):
    n = q.numel()
    if out is None:
        out = torch.empty(n, dtype=dtype, device=q.device)
    if not q.is_cuda:
        out.copy_(dequantize_fp8_block_ref(q, scales, out.dtype))
        return out
    rc = native.lib().dsb200_fp8_block_dequantize(native.ptr(q), native.ptr(scales), native.ptr(out), n, _DT[out.dtype],
                                                  native.stream_ptr())
    native.check(rc)
    return out


def dequant_reduce_fp8_block(q, scales, parts, out=None, dtype=torch.bfloat16, scale_out=1.0):
    """q holds `parts` quantized chunks back to back; returns scale_out * sum of the dequantized chunks."""
    n = q.numel() // parts
    if out is None:
        out = torch.empty(n, dtype=dtype, device=q.device)
    if not q.is_cuda:
        full = dequantize_fp8_block_ref(q, scales, torch.float32).view(parts, n)
        out.copy_((full.sum(0) * scale_out).to(out.dtype))
        return out
    rc = native.lib().dsb200_fp8_block_dequant_reduce(native.ptr(q), native.ptr(scales), native.ptr(out), n, parts, float(scale_out),
                                                      _DT[out.dtype], native.stream_ptr())
    native.check(rc)
    return out


# ----------------------------------------------------------------------------- int8 / int4 (reference API parity)
def int_quantize_ref(x, groups, bits=8, symmetric=True):
    xf = x.float().view(groups, -1)
    if symmetric:
        qmax = 2 ** (bits - 1) - 1
        amax = xf.abs().amax(1)
        scale = torch.where(amax > 0, amax / qmax, torch.ones_like(amax))
        q = torch.clamp(torch.round(xf / scale[:, None]), -qmax - 1, qmax)
        return q, torch.stack([scale, torch.zeros_like(scale)], 1)
    mx, mn = xf.amax(1), xf.amin(1)
    scale = torch.where(mx > mn, (mx - mn) / (2 ** bits - 1), torch.ones_like(mx))
    q = torch.clamp(torch.round((xf - mn[:, None]) / scale[:, None]), 0, 2 ** bits - 1)
    return q, torch.stack([scale, mn], 1)


def int_dequantize_ref(q, params, dtype=torch.float32):
    return (q * params[:, 0:1] + params[:, 1:2]).to(dtype)


def int_quantize(x, groups, bits=8, symmetric=True):
    """Returns (int8 payload [n*bits/8], params [groups, 2] = (scale, offset))."""
    x = x.contiguous().view(-1)
    n = x.numel()
    assert n % groups == 0
    gs = n // groups
    q = torch.empty(n * bits // 8, dtype=torch.int8, device=x.device)
    params = torch.empty(groups, 2, dtype=torch.float32, device=x.device)
    rc = native.lib().dsb200_int_quantize(native.ptr(x), native.ptr(q), native.ptr(params), groups, gs, bits, int(symmetric),
                                          _DT[x.dtype], native.stream_ptr())
    native.check(rc)
    return q, params


def int_dequantize(q, params, groups, bits=8, symmetric=True, dtype=torch.float16):
    n = q.numel() * 8 // bits
    out = torch.empty(n, dtype=dtype, device=q.device)
    rc = native.lib().dsb200_int_dequantize(native.ptr(q), native.ptr(params), native.ptr(out), groups, n // groups, bits,
                                            int(symmetric), _DT[dtype], native.stream_ptr())
    native.check(rc)
    return out


def fake_quantize_(x, groups, bits, symmetric=True, stochastic=False, seed=0):
    """In-place quantize-dequantize (MoQ / QAT).  parity: ds_quantize / ds_sr_quantize (+_asym)."""
    assert x.is_contiguous()
    n = x.numel()
    if not x.is_cuda:
        xf = x.float().view(groups, -1)
        levels = 2 ** bits
        if symmetric:
            amax = xf.abs().amax(1, keepdim=True)
            scale = torch.where(amax > 0, 2 * amax / levels, torch.ones_like(amax))
            q = torch.clamp(torch.round(xf / scale), -levels / 2, levels / 2 - 1) * scale
        else:
            mx, mn = xf.amax(1, keepdim=True), xf.amin(1, keepdim=True)
            scale = torch.where(mx > mn, (mx - mn) / (levels - 1), torch.ones_like(mx))
            q = torch.clamp(torch.round((xf - mn) / scale), 0, levels - 1) * scale + mn
        x.copy_(q.view_as(x).to(x.dtype))
        return x
    rc = native.lib().dsb200_fake_quantize(native.ptr(x), groups, n // groups, bits, int(symmetric), int(stochastic), int(seed),
                                           _DT[x.dtype], native.stream_ptr())
    native.check(rc)
    return x


class Quantizer:
    """Object API of the reference's inference quantizer (`ds_quantizer`)."""

    def __init__(self, q_bits=8, group_size=None, groups=None, symmetric=True):
        self.q_bits, self.group_size, self.groups, self.symmetric = q_bits, group_size, groups, symmetric

    def _groups(self, x):
        return self.groups if self.groups is not None else x.numel() // (self.group_size or x.shape[-1])

    def quantize(self, x):
        return int_quantize(x, self._groups(x), self.q_bits, self.symmetric)

    def dequantize(self, q, params, dtype=torch.float16):
        return int_dequantize(q, params, params.shape[0], self.q_bits, self.symmetric, dtype)


def ds_quantizer(input, groups=1, bit_num=8, sr=False, asym=False):
    return fake_quantize_(input, groups, bit_num, symmetric=not asym, stochastic=sr)
