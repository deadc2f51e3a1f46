"""Vocabulary-parallel unembedding: each rank scores a contiguous range of the vocabulary and the logits are all-gathered
(parity: inference/v2/model_implementations/sharding/unembed.py:12-43)."""
import torch

from .types import ShardingType
from .utils import get_shard_endpoints, shard_param


def shard_unembed_param(param: torch.Tensor, shard_rank: int, num_shards: int) -> torch.Tensor:
    """[vocab, hidden] (or a [vocab] bias) -> this rank's vocabulary range (granularity 1: vocab sizes are arbitrary)."""
    return shard_param(param, ShardingType.OUTER_DIMENSION, shard_rank, num_shards, granularity=1)


def sharded_unembed_dim(vocab_size: int, shard_rank: int, num_shards: int) -> int:
    lo, hi = get_shard_endpoints(vocab_size, shard_rank, num_shards, granularity=1)
    return hi - lo
