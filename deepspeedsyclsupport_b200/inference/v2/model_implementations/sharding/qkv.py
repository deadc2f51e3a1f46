"""Fused QKV projection, laid out [all Q heads; all K heads; all V heads] along the output dimension (parity:
inference/v2/model_implementations/sharding/qkv.py:14-166)."""
from typing import Optional

import torch

from .attn import _check_gqa
from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def _is_mha(n_heads_q, n_heads_kv):
    if n_heads_kv is not None and n_heads_q is None:
        raise ValueError("n_heads_kv should not be passed without n_heads_q")
    return n_heads_q is None or n_heads_kv is None or n_heads_q == n_heads_kv


def shard_qkv_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int, n_heads_q: Optional[int] = None,
                    n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    """Weight [(hq + 2 hkv) * d, in] or bias [(hq + 2 hkv) * d] -> this rank's [local Q; local K; local V]."""
    mha = _is_mha(n_heads_q, n_heads_kv)
    if param is None:
        return None
    if n_heads_q is None:
        if param.shape[0] // 3 % head_size != 0:
            raise ValueError("MHA param shape is not correct")
        n_heads_q = param.shape[0] // head_size // 3
    if n_heads_q < num_shards:
        raise ValueError("There must be at least as many query heads as there are shards.")
    if mha:
        return shard_param(param, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, num_concatenated_matrices=3, granularity=head_size)
    _check_gqa(n_heads_q, n_heads_kv, num_shards)
    if param.shape[0] != head_size * (n_heads_q + 2 * n_heads_kv):
        raise ValueError("GQA param shape is not correct")
    q, kv = param[:head_size * n_heads_q], param[head_size * n_heads_q:]
    if n_heads_kv >= num_shards:
        q = shard_param(q, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, granularity=head_size)
        kv = shard_param(kv, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, num_concatenated_matrices=2, granularity=head_size)
        return torch.cat([q, kv], dim=0)
    # one KV head serves several ranks: each takes that head and its part of the head's query group
    ranks_per_group = num_shards // n_heads_kv
    kv_head, sub_rank = divmod(shard_rank, ranks_per_group)
    group = n_heads_q // n_heads_kv
    k = kv[kv_head * head_size:(kv_head + 1) * head_size]
    v = kv[(n_heads_kv + kv_head) * head_size:(n_heads_kv + kv_head + 1) * head_size]
    q_group = q[group * kv_head * head_size:group * (kv_head + 1) * head_size]
    q_local = shard_param(q_group, ShardingType.OUTER_DIMENSION, sub_rank, ranks_per_group, granularity=head_size)
    return torch.cat([q_local, k, v], dim=0)


def qkv_out_features(in_features: int, shard_rank: int, num_shards: int, head_size: int, n_heads_q: Optional[int] = None,
                     n_heads_kv: Optional[int] = None) -> int:
    """Rows of this rank's fused QKV weight."""
    mha = _is_mha(n_heads_q, n_heads_kv)
    if n_heads_q is not None and in_features != head_size * n_heads_q:
        raise ValueError("in_features is not consistent with n_heads_q and head_size")
    if mha:
        lo, hi = get_shard_endpoints(in_features, shard_rank, num_shards, granularity=head_size)
        return (hi - lo) * 3
    _check_gqa(n_heads_q, n_heads_kv, num_shards)
    group = n_heads_q // n_heads_kv
    if n_heads_kv >= num_shards:
        return (n_heads_kv // num_shards) * head_size * (2 + group)
    ranks_per_group = num_shards // n_heads_kv
    lo, hi = get_shard_endpoints(group * head_size, shard_rank % ranks_per_group, ranks_per_group, granularity=head_size)
    return (hi - lo) + 2 * head_size
