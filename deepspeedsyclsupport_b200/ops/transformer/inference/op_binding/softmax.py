"""(parity: op_binding/softmax.py)  Masked softmax over attention scores [B, H, S, T]: additive mask, ALiBi, causal triangle,
local window, per-layer scale."""
import torch

from .base import BaseOp


class SoftmaxOp(BaseOp):

    def forward(self, attn_scores: torch.Tensor, attn_mask: torch.Tensor, alibi: torch.Tensor, triangular: bool, recompute: bool = False,
                local_attention: bool = False, window_size: int = 1, async_op: bool = False, layer_scale: float = 1.0, head_offset: int = 0):
        s = attn_scores.float() * layer_scale
        B, H, S, T = s.shape
        if alibi is not None and alibi.numel() > 1:
            s = s + alibi.reshape(-1, H, 1, alibi.shape[-1])[..., head_offset:head_offset + H, :, :T].float() if alibi.dim() == 4 else \
                s + alibi.reshape(B, -1, 1, alibi.shape[-1])[:, head_offset:head_offset + H, :, :T].float()
        if attn_mask is not None and attn_mask.numel() > 1:
            m = attn_mask.float()
            while m.dim() < 4:
                m = m[:, None]
            s = s + m
        if triangular or local_attention:
            i = torch.arange(T - S, T, device=s.device)[:, None]
            j = torch.arange(T, device=s.device)[None, :]
            allowed = j <= i if triangular else torch.ones(S, T, dtype=torch.bool, device=s.device)
            if local_attention:
                allowed = allowed & (j > i - window_size)
            s = s.masked_fill(~allowed, float("-inf"))
        return torch.softmax(s, dim=-1).to(attn_scores.dtype)
