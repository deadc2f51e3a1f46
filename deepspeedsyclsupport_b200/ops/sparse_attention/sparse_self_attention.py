"""SparseSelfAttention: softmax(Q K^T / sqrt(d) restricted to a block layout) V for [batch, heads, seq, head_dim]
inputs.  Parity: deepspeed/ops/sparse_attention/sparse_self_attention.py (same constructor / forward signature); fp16
and bf16 are both accepted here."""
import torch
import torch.nn as nn

from ... import comm as dist
from .matmul import MatMul
from .softmax import Softmax
from .sparsity_config import SparsityConfig


class SparseSelfAttention(nn.Module):

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode="add", attn_mask_mode="mul", max_seq_length=2048):
        super().__init__()
        self.sparsity_config = sparsity_config
        self.register_buffer("master_layout", self.sparsity_config.make_layout(max_seq_length))
        self._need_layout_synchronization = True
        self.ops = {}  # sequence length -> (sdd, dsd, softmax) built for THIS module's layout
        self.key_padding_mask_mode = key_padding_mask_mode
        self.attn_mask_mode = attn_mask_mode

    def get_layout(self, L):
        # random layouts differ between ranks unless rank 0's is broadcast once
        if self._need_layout_synchronization and dist.is_initialized():
            dist.broadcast(self.master_layout, src=0)
            self._need_layout_synchronization = False
        if L % self.sparsity_config.block != 0:
            raise ValueError(f"Sequence Length, {L}, needs to be dividable by Block size {self.sparsity_config.block}!")
        nb = L // self.sparsity_config.block
        return self.master_layout[..., :nb, :nb].cpu()

    def get_ops(self, H, L):
        key = L
        if key not in self.ops:
            layout = self.get_layout(L)
            blk = self.sparsity_config.block
            self.ops[key] = (MatMul(layout, blk, "sdd", trans_a=False, trans_b=True), MatMul(layout, blk, "dsd", trans_a=False, trans_b=False),
                                            Softmax(layout, blk))
        return self.ops[key]

    def transpose_key_for_scores(self, x, L):
        return x.permute(0, 1, 3, 2) if x.size(2) != L else x

    def transpose_mask_for_sparse(self, qtype, x, is_key_padding_mask=False):
        x = x.type(qtype)
        if is_key_padding_mask:
            while x.dim() > 2:
                x = x.squeeze(1)
            return x
        return x.squeeze()

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        assert query.dtype in (torch.half, torch.bfloat16, torch.float32)
        bsz, num_heads, tgt_len, head_dim = query.size()
        key = self.transpose_key_for_scores(key, tgt_len)
        if query.shape != key.shape or key.shape != value.shape:
            raise NotImplementedError("only self-attention is supported for now")
        if key_padding_mask is not None:
            key_padding_mask = self.transpose_mask_for_sparse(query.dtype, key_padding_mask, is_key_padding_mask=True)
        if attn_mask is not None:
            attn_mask = self.transpose_mask_for_sparse(query.dtype, attn_mask)
        sdd, dsd, softmax = self.get_ops(num_heads, tgt_len)
        scores = sdd(query, key)
        probs = softmax(scores, scale=float(head_dim)**-0.5, rpe=rpe, key_padding_mask=key_padding_mask, attn_mask=attn_mask,
                        key_padding_mask_mode=self.key_padding_mask_mode, attn_mask_mode=self.attn_mask_mode)
        return dsd(probs, value)
