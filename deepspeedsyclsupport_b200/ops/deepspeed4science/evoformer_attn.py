"""DS4Sci_EvoformerAttention: attention with up to two additive biases, as used by AlphaFold/OpenFold Evoformer blocks
(MSA row/column attention and triangle attention).

    Q, K, V : [B, N, L, H, D]        bias1 : [B, N, 1, 1, L]  (mask bias)        bias2 : [B, 1, H, L, L]  (pair bias)
    O = softmax(Q K^T / sqrt(D) + bias1 + bias2) V                                  (same layout as Q)

Reference: csrc/deepspeed4science/evoformer_attn (CUTLASS fused MHA with bias gradients),
deepspeed/ops/deepspeed4science/evoformer_attn.py.  On the GPU (bf16 / fp16, head size <= 64) the op runs on the repo's own
kernels (csrc/attn/evoformer.cu): a fused forward that never materialises the L x L scores and saves the row log-sum-exp, and a
two-pass backward (row pass: dQ, dbias1, dbias2; column pass: dK, dV) that recomputes the probabilities from it.  Other inputs
(CPU, fp32, larger heads) take the memory-efficient SDPA formulation with the two biases broadcast-added into one mask."""
import math

import torch
import torch.nn.functional as F

from .. import native

native.register_sigs({"dsb200_evoformer_attn_fwd": "ppppppp iiiii f i p", "dsb200_evoformer_attn_bwd": "pppppppp ppp ppp iiiii f i p"})
_DT = {torch.bfloat16: 0, torch.float16: 1}


def _native_ok(q, k, v, b1, b2):
    if not (q.is_cuda and q.dtype in _DT and q.dim() == 5 and q.shape[-1] <= 64 and k.shape == q.shape and v.shape == q.shape):
        return False
    return all(b is None or b.is_cuda for b in (b1, b2))


class _EvoformerNative(torch.autograd.Function):
    """csrc/attn/evoformer.cu: fused forward (+ LSE), two-pass backward with bias gradients."""

    @staticmethod
    def forward(ctx, q, k, v, bias1, bias2):
        B, N, L, H, D = q.shape
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        b1 = None if bias1 is None else bias1.to(q.dtype).contiguous()
        b2 = None if bias2 is None else bias2.to(q.dtype).contiguous()
        o = torch.empty_like(q)
        lse = torch.empty(B, N, H, L, dtype=torch.float32, device=q.device)
        scale = 1.0 / math.sqrt(D)
        rc = native.lib().dsb200_evoformer_attn_fwd(native.ptr(q), native.ptr(k), native.ptr(v), native.ptr(b1), native.ptr(b2), native.ptr(o),
                                                    native.ptr(lse), B, N, L, H, D, scale, _DT[q.dtype], native.stream_ptr())
        native.check(rc)
        ctx.save_for_backward(q, k, v, o, lse, *(t for t in (b1, b2) if t is not None))
        ctx.has = (b1 is not None, b2 is not None)
        ctx.bias_meta = tuple(None if b is None else (b.dtype, b.shape) for b in (bias1, bias2))
        return o

    @staticmethod
    def backward(ctx, d_o):
        q, k, v, o, lse, *rest = ctx.saved_tensors
        b1 = rest.pop(0) if ctx.has[0] else None
        b2 = rest.pop(0) if ctx.has[1] else None
        B, N, L, H, D = q.shape
        d_o = d_o.contiguous()
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        delta = torch.empty_like(lse)
        want1, want2 = b1 is not None and ctx.needs_input_grad[3], b2 is not None and ctx.needs_input_grad[4]
        db1 = torch.zeros(B, N, L, dtype=torch.float32, device=q.device) if want1 else None
        db2 = torch.zeros(B, H, L, L, dtype=torch.float32, device=q.device) if want2 else None
        rc = native.lib().dsb200_evoformer_attn_bwd(native.ptr(q), native.ptr(k), native.ptr(v), native.ptr(b1), native.ptr(b2), native.ptr(o),
                                                    native.ptr(d_o), native.ptr(lse), native.ptr(dq), native.ptr(dk), native.ptr(dv),
                                                    native.ptr(delta), native.ptr(db1), native.ptr(db2), B, N, L, H, D, 1.0 / math.sqrt(D),
                                                    _DT[q.dtype], native.stream_ptr())
        native.check(rc)
        g1 = db1.to(ctx.bias_meta[0][0]).view(ctx.bias_meta[0][1]) if want1 else None
        g2 = db2.to(ctx.bias_meta[1][0]).view(ctx.bias_meta[1][1]) if want2 else None
        return dq, dk, dv, g1, g2



class EvoformerFusedAttention(torch.autograd.Function):

    @staticmethod
    def forward(ctx, q, k, v, bias1=None, bias2=None):
        """q, k, v: [*, L, H, D]"""
        with torch.enable_grad():
            qd, kd, vd = (t.detach().requires_grad_(t.requires_grad) for t in (q, k, v))
            b1 = bias1.detach().requires_grad_(bias1.requires_grad) if bias1 is not None else None
            b2 = bias2.detach().requires_grad_(bias2.requires_grad) if bias2 is not None else None
            mask = None
            if b1 is not None and b2 is not None:
                mask = b1 + b2
            elif b1 is not None:
                mask = b1
            elif b2 is not None:
                mask = b2
            qh, kh, vh = (t.transpose(-2, -3) for t in (qd, kd, vd))  # [*, H, L, D]
            if mask is not None:
                mask = mask.expand(*qh.shape[:-1], kh.shape[-2]).to(qh.dtype)
            out = F.scaled_dot_product_attention(qh, kh, vh, attn_mask=mask).transpose(-2, -3).contiguous()
        ctx.graph = (out, qd, kd, vd, b1, b2)
        return out.detach()

    @staticmethod
    def backward(ctx, grad_output):
        out, qd, kd, vd, b1, b2 = ctx.graph
        inputs = [t for t in (qd, kd, vd, b1, b2) if t is not None and t.requires_grad]
        grads = list(torch.autograd.grad(out, inputs, grad_output.contiguous(), allow_unused=True)) if inputs else []
        result = []
        for t in (qd, kd, vd, b1, b2):
            result.append(grads.pop(0) if (t is not None and t.requires_grad) else None)
        ctx.graph = None
        return tuple(result)


def DS4Sci_EvoformerAttention(Q, K, V, biases):
    assert len(biases) <= 2
    biases = list(biases) + [None] * (2 - len(biases))
    if biases[0] is not None:
        assert biases[0].shape == (Q.shape[0], Q.shape[1], 1, 1, Q.shape[2]), "bias1 shape is incorrect"
    if biases[1] is not None:
        assert biases[1].shape == (Q.shape[0], 1, Q.shape[3], Q.shape[2], Q.shape[2]), "bias2 shape is incorrect"
    if _native_ok(Q, K, V, biases[0], biases[1]):
        return _EvoformerNative.apply(Q, K, V, biases[0], biases[1])
    return EvoformerFusedAttention.apply(Q, K, V, biases[0], biases[1])
