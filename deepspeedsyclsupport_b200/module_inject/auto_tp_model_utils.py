"""Head-sliced position biases for HuggingFace models run under AutoTP: BLOOM / MPT / Baichuan build their ALiBi (or
attention-bias) tensors for all heads, while a tensor-parallel rank owns a contiguous slice of the heads.  Each function below
replaces the model's builder and returns only this rank's heads.
(reference: deepspeed/module_inject/auto_tp_model_utils.py)"""
import math

import torch

from .. import comm as dist
from .tp_shard import get_shard_size, get_shard_size_list


def _my_heads(num_heads):
    """(first head, number of heads) owned by this rank; the whole range outside a distributed run."""
    if not dist.is_initialized():
        return 0, num_heads
    world, rank = dist.get_world_size(), dist.get_rank()
    return sum(get_shard_size_list(num_heads, world)[:rank]), get_shard_size(num_heads, world)


def alibi_slopes(num_heads, device=None):
    """Geometric slope sequence of the ALiBi paper, including the interleaved extension for non-power-of-two head counts."""
    def pow2_slopes(n):
        start = 2.0**(-(2.0**-(math.log2(n) - 3)))
        return torch.tensor([start**(i + 1) for i in range(n)], dtype=torch.float32, device=device)

    n = 2**math.floor(math.log2(num_heads))
    slopes = pow2_slopes(n)
    if n != num_heads:
        extra = pow2_slopes(2 * n)[0::2][:num_heads - n]
        slopes = torch.cat([slopes, extra])
    return slopes


def build_bloom_alibi_tensor(attention_mask, num_heads, dtype):
    """[batch * local_heads, 1, seq]: slope_h * (position of the token among the attended ones)."""
    batch, seq = attention_mask.shape
    first, count = _my_heads(num_heads)
    slopes = alibi_slopes(num_heads, attention_mask.device)[first:first + count]
    positions = ((attention_mask.cumsum(dim=-1) - 1) * attention_mask)[:, None, :]
    alibi = slopes[None, :, None] * positions  # [batch, local_heads, seq]
    return alibi.reshape(batch * count, 1, seq).to(dtype)


def get_alibi_mask(self, tensor, seq_length_with_past):
    """Baichuan: the model's own mask [heads, S, S], restricted to this rank's heads at inference time."""
    mask = self.get_alibi_mask_orig(tensor, seq_length_with_past)
    if not self.training and dist.is_initialized():
        first, count = _my_heads(self.n_head)
        mask = mask[first:first + count, :seq_length_with_past, :seq_length_with_past]
    return mask


def build_mpt_atten_bias_tensor(self, device, dtype, attention_mask=None, prefix_mask=None, sequence_id=None):
    attn_bias, attention_mask = self._attn_bias_orig(device, dtype, attention_mask=attention_mask, prefix_mask=prefix_mask,
                                                     sequence_id=sequence_id)
    if dist.is_initialized():
        first, count = _my_heads(self.config.n_heads)
        attn_bias = attn_bias[:, first:first + count, :, :]
    return attn_bias, attention_mask


def build_mpt_alibi_tensor(self, num_heads, sequence_length, alibi_bias_max=8, device=None):
    alibi = self.build_mpt_alibi_tensor_orig(num_heads, sequence_length, alibi_bias_max, device)
    if dist.is_initialized():
        first, count = _my_heads(num_heads)
        alibi = alibi[first:first + count, :, :]
    return alibi
