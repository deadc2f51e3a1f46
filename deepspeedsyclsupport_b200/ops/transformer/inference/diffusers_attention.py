"""Attention of the diffusers UNet / VAE blocks (parity: deepspeed/ops/transformer/inference/diffusers_attention.py:34-196):
self-attention over image tokens with a fused QKV projection, or cross-attention whose keys / values come from the text-encoder
context (separate q / k / v weights because the widths differ).  No mask, no KV cache; the softmax(QK^T)V core is the fused
SDPA kernel, the projections the tcgen05 GEMM for bf16."""
import math

import torch
import torch.nn as nn

from ....accelerator import get_accelerator
from .op_binding import LinearOp, VectorMatMulOp


class DeepSpeedDiffusersAttention(nn.Module):
    """`config`: a DeepSpeedInferenceConfig (hidden_size, heads, dtype, mp_size).  Parameters are created empty and filled by the
    injection policy (UNetPolicy.attention)."""
    layer_id = 0

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.config.layer_id = DeepSpeedDiffusersAttention.layer_id
        DeepSpeedDiffusersAttention.layer_id += 1
        dev = get_accelerator().current_device_name() if config.dtype in (torch.float16, torch.bfloat16) else "cpu"
        dt = torch.half if config.dtype == torch.int8 else config.dtype
        h, mp = config.hidden_size, config.mp_size
        mk = lambda *s: nn.Parameter(torch.empty(*s, dtype=dt, device=dev), requires_grad=False)
        self.attn_qkvw = mk(3 * h // mp, h)  # self-attention: fused [Q; K; V]
        self.attn_qw, self.attn_kw, self.attn_vw = mk(h // mp, h), mk(h // mp, h), mk(h // mp, h)  # cross-attention (k / v re-shaped on fill)
        self.attn_qkvb = mk(3 * h // mp)
        self.attn_ow, self.attn_ob = mk(h, h // mp), mk(h)
        self.do_out_bias = True
        self.num_attention_heads_per_partition = config.heads // mp
        self.hidden_size_per_partition = h // mp
        self.hidden_size_per_attention_head = h // config.heads
        self.norm_factor = math.sqrt(math.sqrt(h // config.heads))
        self.scale = 1.0 / math.sqrt(h // config.heads)
        self.linear_func = LinearOp(config)
        self.vector_matmul_func = VectorMatMulOp(config)

    def _heads(self, x):
        B, T, _ = x.shape
        return x.view(B, T, self.num_attention_heads_per_partition, -1).transpose(1, 2)

    def forward(self, input, context=None, input_mask=None):
        """input [B, tokens, hidden]; context [B, text tokens, context width] for cross-attention."""
        if self.config.layer_id == 0 and hasattr(self, "_warm") is False:
            self._warm = True
        if context is None:
            qkv = self.linear_func(input, self.attn_qkvw, self.attn_qkvb if self.attn_qkvb is not None else None, add_bias=self.attn_qkvb is not None)
            q, k, v = qkv.chunk(3, dim=-1)
        else:
            q = self.vector_matmul_func(input, self.attn_qw)
            k = self.vector_matmul_func(context, self.attn_kw)
            v = self.vector_matmul_func(context, self.attn_vw)
        ctx = torch.nn.functional.scaled_dot_product_attention(self._heads(q), self._heads(k), self._heads(v), scale=self.scale)
        ctx = ctx.transpose(1, 2).reshape(input.shape[0], input.shape[1], -1)
        out = self.vector_matmul_func(ctx, self.attn_ow)
        if self.config.mp_size > 1:
            from .... import comm as dist
            dist.all_reduce(out)
        return out + self.attn_ob if self.do_out_bias else out
