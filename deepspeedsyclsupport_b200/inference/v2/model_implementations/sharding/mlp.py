"""MLP projections, dense or stacked per expert ([E, out, in]) (parity: inference/v2/model_implementations/sharding/mlp.py:14-77)."""
from typing import Optional

import torch

from .types import DEFAULT_SHARD_GRANULARITY, ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_mlp_1_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, gated: bool = False,
                      is_moe: bool = False) -> Optional[torch.Tensor]:
    """Up projection (column-parallel).  gated: the weight is [gate; up] stacked along the output dimension and both halves are
    cut at the same place, so act(gate) * up stays local."""
    return shard_param(param, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, num_concatenated_matrices=2 if gated else 1,
                       bias_dims=2 if is_moe else 1)


def shard_mlp_2_param(param: Optional[torch.Tensor], shard_rank: int, num_shards: int, is_moe: bool = False) -> Optional[torch.Tensor]:
    """Down projection (row-parallel).  Its bias is applied once, after the all-reduce: rank 0 keeps it."""
    if param is None:
        return None
    if param.ndim == (2 if is_moe else 1):
        return param if shard_rank == 0 else None
    return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards)


def sharded_intermediate_dim(intermediate_size: int, num_shards: int, shard_rank: int) -> int:
    lo, hi = get_shard_endpoints(intermediate_size, shard_rank, num_shards, granularity=DEFAULT_SHARD_GRANULARITY)
    return hi - lo
