"""(parity: inference/v2/model_implementations/sharding/types.py)"""
from enum import Enum

DEFAULT_SHARD_GRANULARITY = 32


class ShardingType(Enum):
    OUTER_DIMENSION = 0  # column-parallel linear: output features (rows of an nn.Linear weight, entries of its bias)
    INNER_DIMENSION = 1  # row-parallel linear: input features (columns of the weight); the bias is not sharded
