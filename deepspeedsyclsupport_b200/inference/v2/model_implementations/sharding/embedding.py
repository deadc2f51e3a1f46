"""Token / position embedding tables are cut along the hidden dimension; ranks all-gather the looked-up rows (parity:
inference/v2/model_implementations/sharding/embedding.py:12-34).  The B200 engine keeps the table whole on every rank when it
fits (one gather launch, no collective), so this rule is only applied when the caller asks for it."""
import torch

from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_embedding_param(param: torch.Tensor, shard_rank: int, num_shards: int) -> torch.Tensor:
    return shard_param(param, ShardingType.INNER_DIMENSION, shard_rank, num_shards)


def sharded_embedding_dim(embedding_size: int, shard_rank: int, num_shards: int) -> int:
    lo, hi = get_shard_endpoints(embedding_size, shard_rank, num_shards)
    return hi - lo
