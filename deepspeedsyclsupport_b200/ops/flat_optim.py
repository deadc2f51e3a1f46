"""Flat-shard optimizer steps shared by the ZeRO optimizers and the list-API optimizer classes.

One call updates a whole contiguous fp32 shard (master weights + state) and writes the 16-bit parameter copy in
the same pass.  CUDA tensors go to csrc/optim/fused_optim.cu, host tensors to csrc/host/cpu_optim.cpp (AVX-512),
with a pure-PyTorch fallback used only as the numerical oracle.
"""
import ctypes
import math

import torch

from . import native

_GT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}
_LP = {None: 0, torch.bfloat16: 1, torch.float16: 2}


def _host_sigs():
    l = native.host_lib()
    if getattr(l, "_sigs_done", False):
        return l
    P, I, L, F = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float
    l.dsbh_adam_step.argtypes = [P, P, P, P, I, P, I, L, F, F, F, F, F, I, I, I, F]
    l.dsbh_adagrad_step.argtypes = [P, P, P, I, P, I, L, F, F, F, F]
    l.dsbh_lion_step.argtypes = [P, P, P, I, P, I, L, F, F, F, F, F]
    l.dsbh_cast_f32_lp.argtypes = [P, P, I, L]
    l.dsbh_sumsq.argtypes = [P, I, L]
    l.dsbh_sumsq.restype = ctypes.c_double
    l._sigs_done = True
    return l


def adam_ref(p, m, v, g, lr, beta1, beta2, eps, weight_decay, adamw_mode, bias_correction, step, grad_scale=1.0):
    """fp32 oracle with the exact update order of the kernels."""
    g = g.float() * grad_scale
    if not adamw_mode:
        g = g + weight_decay * p
    m.mul_(beta1).add_(g, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    denom = v.sqrt() / math.sqrt(bc2) + eps
    upd = (m / bc1) / denom
    if adamw_mode:
        upd = upd + weight_decay * p
    p.add_(upd, alpha=-lr)


def adam_step(p, m, v, g, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, adamw_mode=True,
              bias_correction=True, step=1, lp=None, inv_scale=None, skip=None, grad_scale=1.0):
    """In-place Adam/AdamW on flat fp32 `p, m, v` with gradient `g` (fp32/bf16/fp16).
    lp: optional 16-bit tensor receiving the updated parameters.  inv_scale/skip: CUDA scalars (float / int32)."""
    n = p.numel()
    if n == 0:
        return
    if p.is_cuda:
        rc = native.lib().dsb200_fused_adam_flat(native.ptr(p), native.ptr(m), native.ptr(v), native.ptr(g), _GT[g.dtype],
                                                 native.ptr(lp), _LP[lp.dtype if lp is not None else None], n, lr, beta1, beta2,
                                                 eps, weight_decay, int(adamw_mode), int(bias_correction), int(step),
                                                 native.ptr(inv_scale), native.ptr(skip), native.stream_ptr())
        native.check(rc)
        return
    gs = grad_scale * (float(inv_scale) if inv_scale is not None else 1.0)
    if skip is not None and int(skip) != 0:
        return
    try:
        l = _host_sigs()
    except native.NativeError:
        l = None
    if l is not None and p.is_contiguous() and g.is_contiguous():
        l.dsbh_adam_step(p.data_ptr(), m.data_ptr(), v.data_ptr(), g.data_ptr(), _GT[g.dtype],
                         lp.data_ptr() if lp is not None else None, _LP[lp.dtype if lp is not None else None], n, lr, beta1,
                         beta2, eps, weight_decay, int(adamw_mode), int(bias_correction), int(step), gs)
    else:
        adam_ref(p, m, v, g, lr, beta1, beta2, eps, weight_decay, adamw_mode, bias_correction, step, gs)
        if lp is not None:
            lp.copy_(p)


def lion_step(p, m, g, *, lr, beta1=0.9, beta2=0.99, weight_decay=0.0, lp=None, inv_scale=None, skip=None, grad_scale=1.0):
    n = p.numel()
    if n == 0:
        return
    if p.is_cuda:
        rc = native.lib().dsb200_fused_lion_flat(native.ptr(p), native.ptr(m), native.ptr(g), _GT[g.dtype], native.ptr(lp),
                                                 _LP[lp.dtype if lp is not None else None], n, lr, beta1, beta2, weight_decay,
                                                 native.ptr(inv_scale), native.ptr(skip), native.stream_ptr())
        native.check(rc)
        return
    gs = grad_scale * (float(inv_scale) if inv_scale is not None else 1.0)
    if skip is not None and int(skip) != 0:
        return
    l = _host_sigs()
    l.dsbh_lion_step(p.data_ptr(), m.data_ptr(), g.data_ptr(), _GT[g.dtype], lp.data_ptr() if lp is not None else None,
                     _LP[lp.dtype if lp is not None else None], n, lr, beta1, beta2, weight_decay, gs)


def adagrad_step(p, h, g, *, lr, eps=1e-10, weight_decay=0.0, lp=None, inv_scale=None, skip=None, grad_scale=1.0):
    n = p.numel()
    if n == 0:
        return
    if p.is_cuda:
        rc = native.lib().dsb200_fused_adagrad_flat(native.ptr(p), native.ptr(h), native.ptr(g), _GT[g.dtype], native.ptr(lp),
                                                    _LP[lp.dtype if lp is not None else None], n, lr, eps, weight_decay,
                                                    native.ptr(inv_scale), native.ptr(skip), native.stream_ptr())
        native.check(rc)
        return
    gs = grad_scale * (float(inv_scale) if inv_scale is not None else 1.0)
    if skip is not None and int(skip) != 0:
        return
    l = _host_sigs()
    l.dsbh_adagrad_step(p.data_ptr(), h.data_ptr(), g.data_ptr(), _GT[g.dtype], lp.data_ptr() if lp is not None else None,
                        _LP[lp.dtype if lp is not None else None], n, lr, eps, weight_decay, gs)


def lamb_step(p, m, v, g, *, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, bias_correction=True, step=1,
              max_coeff=10.0, min_coeff=0.01, lp=None, inv_scale=None, skip=None):
    """LAMB on ONE tensor (trust ratio is per tensor).  Returns the trust ratio tensor (CUDA) or float (CPU)."""
    n = p.numel()
    if p.is_cuda:
        lib = native.lib()
        u = torch.empty(n, dtype=torch.float32, device=p.device)
        scratch = torch.empty(lib.dsb200_fused_lamb_scratch_floats(n), dtype=torch.float32, device=p.device)
        coeff = torch.empty(1, dtype=torch.float32, device=p.device)
        rc = lib.dsb200_fused_lamb_flat(native.ptr(p), native.ptr(m), native.ptr(v), native.ptr(g), _GT[g.dtype], native.ptr(lp),
                                        _LP[lp.dtype if lp is not None else None], native.ptr(u), native.ptr(scratch), n, lr,
                                        beta1, beta2, eps, weight_decay, int(bias_correction), int(step), max_coeff, min_coeff,
                                        native.ptr(inv_scale), native.ptr(skip), native.ptr(coeff), native.stream_ptr())
        native.check(rc, 2)
        return coeff
    gg = g.float() * (float(inv_scale) if inv_scale is not None else 1.0)
    m.mul_(beta1).add_(gg, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(gg, gg, value=1 - beta2)
    bc1 = 1 - beta1 ** step if bias_correction else 1.0
    bc2 = 1 - beta2 ** step if bias_correction else 1.0
    upd = (m / bc1) / ((v / bc2).sqrt() + eps) + weight_decay * p
    wn, un = p.norm().item(), upd.norm().item()
    ratio = min(max(wn / un, min_coeff), max_coeff) if wn > 0 and un > 0 else 1.0
    p.add_(upd, alpha=-lr * ratio)
    if lp is not None:
        lp.copy_(p)
    return ratio
