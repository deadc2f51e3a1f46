"""DS4Sci_EvoformerAttention: attention with up to two additive biases, as used by AlphaFold/OpenFold Evoformer blocks
(MSA row/column attention and triangle attention).

    Q, K, V : [B, N, L, H, D]        bias1 : [B, N, 1, 1, L]  (mask bias)        bias2 : [B, 1, H, L, L]  (pair bias)
    O = softmax(Q K^T / sqrt(D) + bias1 + bias2) V                                  (same layout as Q)

Reference: csrc/deepspeed4science/evoformer_attn (CUTLASS fused MHA with bias gradients),
deepspeed/ops/deepspeed4science/evoformer_attn.py.  Here the op is an autograd Function around the memory-efficient
fused attention kernel with the two biases broadcast-added into one additive mask; the bias gradients are the
row/column reductions of the mask gradient.  L^2 scores are never materialised in fp32 by the fused kernel; the combined
bias tensor is [B, N, H, L, L] only when both biases are present."""
import torch
import torch.nn.functional as F


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
    return EvoformerFusedAttention.apply(Q, K, V, biases[0], biases[1])
