"""Slicing arithmetic shared by every tensor-parallel sharding rule (parity: inference/v2/model_implementations/sharding/
utils.py:14-104).

A dimension is cut in units of `granularity` (head size for attention, 32 elsewhere so GEMM tiles stay aligned); when the units
do not divide evenly the first `units % num_shards` ranks take one extra unit."""
from typing import Optional, Tuple

import torch

from .types import DEFAULT_SHARD_GRANULARITY, ShardingType


def get_shard_endpoints(dim_size: int, shard_rank: int, num_shards: int, granularity: int = DEFAULT_SHARD_GRANULARITY) -> Tuple[int, int]:
    assert dim_size % granularity == 0, "Dimension size must be divisible by granularity"
    units = dim_size // granularity
    base, extra = divmod(units, num_shards)
    start = shard_rank * base + min(shard_rank, extra)
    stop = start + base + (1 if shard_rank < extra else 0)
    return start * granularity, stop * granularity


def _slice_each(param: torch.Tensor, dim: int, pieces: int, shard_rank: int, num_shards: int, granularity: int) -> torch.Tensor:
    """`param` is `pieces` matrices concatenated along `dim` (fused QKV: 3, gate/up: 2): take this rank's range of each."""
    lo, hi = get_shard_endpoints(param.size(dim) // pieces, shard_rank, num_shards, granularity)
    parts = [m.narrow(dim, lo, hi - lo) for m in torch.chunk(param, pieces, dim=dim)]
    return parts[0] if pieces == 1 else torch.cat(parts, dim=dim)


def shard_param(param: Optional[torch.Tensor], shard_mode: ShardingType, shard_rank: int, num_shards: int,
                num_concatenated_matrices: int = 1, granularity: int = DEFAULT_SHARD_GRANULARITY, bias_dims: int = 1) -> Optional[torch.Tensor]:
    """This rank's slice of a weight ([..., out, in]) or bias ([..., out]; a tensor with `bias_dims` dimensions is a bias --
    2 for stacked MoE biases)."""
    assert shard_rank < num_shards, "Shard rank must be less than num_shards"
    if param is None:
        return None
    if num_shards == 1:
        return param
    if shard_mode == ShardingType.OUTER_DIMENSION:
        dim = -1 if param.ndim == bias_dims else -2
        return _slice_each(param, dim, num_concatenated_matrices, shard_rank, num_shards, granularity)
    if shard_mode == ShardingType.INNER_DIMENSION:
        return _slice_each(param, -1, num_concatenated_matrices, shard_rank, num_shards, granularity)
    raise ValueError(f"unknown sharding mode {shard_mode}")
