"""Attention output projection: row-parallel over the heads each rank computed (parity: inference/v2/model_implementations/
sharding/attn_out.py:14-123)."""
from typing import Optional

import torch

from .attn import _check_gqa
from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_attn_out_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, head_size: int, n_heads_q: Optional[int] = None,
                         n_heads_kv: Optional[int] = None) -> Optional[torch.Tensor]:
    """Weight [out, hq * d] -> columns of this rank's query heads.  The bias is added after the all-reduce, so rank 0 keeps it
    and the others get None."""
    if param is None:
        return None
    if param.ndim == 1:
        return param if shard_rank == 0 else None
    if n_heads_kv is None or n_heads_kv == n_heads_q:
        return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards, granularity=head_size)
    _check_gqa(n_heads_q, n_heads_kv, num_shards)
    if n_heads_kv >= num_shards:
        group_cols = (n_heads_q // n_heads_kv) * head_size  # whole query groups travel with their KV head
        return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards, granularity=group_cols)
    ranks_per_group = num_shards // n_heads_kv
    kv_head, sub_rank = divmod(shard_rank, ranks_per_group)
    group_cols = (n_heads_q // n_heads_kv) * head_size
    chunk = param[..., kv_head * group_cols:(kv_head + 1) * group_cols]
    return shard_param(chunk, ShardingType.INNER_DIMENSION, sub_rank, ranks_per_group, granularity=head_size)


def attn_out_in_features(out_features: int, shard_rank: int, num_shards: int, head_size: int, n_heads_q: Optional[int] = None,
                         n_heads_kv: Optional[int] = None) -> int:
    """Columns of this rank's attention-output weight (= local query heads * head size)."""
    if n_heads_kv is None or n_heads_kv == n_heads_q:
        # square projection in every MHA model: the head dimension equals out_features unless the head count says otherwise
        dim = n_heads_q * head_size if n_heads_q is not None else out_features
        lo, hi = get_shard_endpoints(dim, shard_rank, num_shards, granularity=head_size)
        return hi - lo
    from .attn import get_local_heads
    return get_local_heads(shard_rank, num_shards, n_heads_q, n_heads_kv)[0] * head_size
