"""Hand-written sm_100a flash attention (csrc/attn) on the packed QKV projection, with autograd."""
import math

import torch

from . import native
from . import recompute as _rc
from ..profiling.flops_profiler import profiler as _prof

native.register_sigs({
    "dsb200_flash_attn_fwd": "ppp iiiii i f p",
    "dsb200_flash_attn_bwd": "pppp ppp iiiii i f p",
})


def flash_fwd(qkv, B, S, hq, hkv, d, causal=True, scale=None, need_lse=True):
    assert qkv.dtype == torch.bfloat16 and qkv.is_contiguous()
    scale = scale if scale is not None else 1.0 / math.sqrt(d)
    out = torch.empty(B * S, hq * d, dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty(B, hq, S, dtype=torch.float32, device=qkv.device) if need_lse else None
    rc = native.lib().dsb200_flash_attn_fwd(native.ptr(qkv), native.ptr(out), native.ptr(lse), B, S, hq, hkv, d, int(causal),
                                            float(scale), native.stream_ptr())
    native.check(rc)
    if _prof._ACTIVE:
        macs = 2 * B * hq * S * S * d // (2 if causal else 1)
        _prof.record_flops(2 * macs, macs, "dsb200_flash_attn_fwd")
    return out, lse


def flash_bwd(dout, qkv, out, lse, B, S, hq, hkv, d, causal=True, scale=None):
    scale = scale if scale is not None else 1.0 / math.sqrt(d)
    dqkv = torch.empty_like(qkv)
    # scratch: fp32 dQ accumulator [B*S, hq*d] (TMA reduce-add target) + delta [B, hq, S]
    dq_acc = torch.zeros(B * S, hq * d, dtype=torch.float32, device=qkv.device)
    delta = torch.empty(B, hq, S, dtype=torch.float32, device=qkv.device)
    rc = native.lib().dsb200_flash_attn_bwd(native.ptr(dout), native.ptr(qkv), native.ptr(out), native.ptr(lse), native.ptr(dqkv),
                                            native.ptr(dq_acc), native.ptr(delta), B, S, hq, hkv, d, int(causal), float(scale),
                                            native.stream_ptr())
    native.check(rc, 3)
    return dqkv


# Under activation checkpointing the attention core's (out, lse) can be kept across the region and replayed during
# recomputation (ops/recompute.py); `attention_recompute_contexts` is the "attention" policy of that module.
def attention_recompute_contexts():
    return _rc.contexts("attention")


class _FlashPackedFn(torch.autograd.Function):

    @staticmethod
    def forward(ctx, qkv, B, S, hq, hkv, d, causal, scale):
        kept = _rc.replayed(_rc.ATTN)
        if kept is not None:
            out, lse = kept
        else:
            out, lse = flash_fwd(qkv, B, S, hq, hkv, d, causal, scale)
            _rc.record(_rc.ATTN, (out, lse))
        ctx.save_for_backward(qkv, out, lse)
        ctx.cfg = (B, S, hq, hkv, d, causal, scale)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse = ctx.saved_tensors
        B, S, hq, hkv, d, causal, scale = ctx.cfg
        if not dout.is_contiguous():
            dout = dout.contiguous()
        dqkv = flash_bwd(dout, qkv, out, lse, B, S, hq, hkv, d, causal, scale)
        return dqkv, None, None, None, None, None, None, None


def flash_attention_packed(qkv, B, S, hq, hkv, d, causal=True, scale=None):
    qkv = qkv if qkv.is_contiguous() else qkv.contiguous()
    if not (torch.is_grad_enabled() and qkv.requires_grad):
        return flash_fwd(qkv, B, S, hq, hkv, d, causal, scale, need_lse=False)[0]
    return _FlashPackedFn.apply(qkv, B, S, hq, hkv, d, causal, scale)
