"""Causal multi-head / grouped-query attention on a packed QKV projection.

`attention(qkv, B, S, hq, hkv, d)` consumes the [B*S, (hq + 2*hkv) * d] output of the fused QKV GEMM directly
(no head-layout transposes: the kernel reads strided head slices) and returns [B*S, hq*d].
On CUDA the hand-written sm_100a flash-attention kernels are used when built (csrc/attn); the PyTorch SDPA form
below is the numerical oracle and the CPU path.
"""
import math

import torch
import torch.nn.functional as F

from . import native

_own_kernel = None


def attention_ref(qkv, B, S, hq, hkv, d, causal=True, scale=None):
    q, k, v = qkv.view(B, S, hq + 2 * hkv, d).split([hq, hkv, hkv], dim=2)
    q, k, v = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    if hkv != hq:
        k = k.repeat_interleave(hq // hkv, dim=1)
        v = v.repeat_interleave(hq // hkv, dim=1)
    o = F.scaled_dot_product_attention(q, k, v, is_causal=causal, scale=scale)
    return o.transpose(1, 2).reshape(B * S, hq * d)


def _have_own():
    global _own_kernel
    if _own_kernel is None:
        # a GPU box without the native library is a broken install: fail loudly instead of silently running a library path
        _own_kernel = hasattr(native.lib(), "dsb200_flash_attn_fwd")
    return _own_kernel


def attention(qkv, B, S, hq, hkv, d, causal=True, scale=None):
    if qkv.is_cuda and _have_own() and qkv.dtype == torch.bfloat16 and d in (64, 128):
        from .flash_attention import flash_attention_packed
        return flash_attention_packed(qkv, B, S, hq, hkv, d, causal, scale)
    return attention_ref(qkv, B, S, hq, hkv, d, causal, scale)
