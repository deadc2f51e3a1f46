"""State-manager / KV-cache configuration (parity: inference/v2/ragged/manager_configs.py:137-183)."""
from enum import Enum
from typing import Optional, Tuple

from pydantic import Field, model_validator

from ....runtime.config_utils import DeepSpeedConfigModel


class KVCacheType(Enum):
    DENSE = "dense"
    LOCAL = "local"


class AllocationMode(Enum):
    RESERVE = "reserve"
    ALLOCATE = "allocate"


class KVCacheConfig(DeepSpeedConfigModel):
    block_size: int = 128
    num_allocation_groups: int = 1
    cache_shape: Tuple[int, int, int]  # (num_layers, num_kv_heads, head_size)
    cache_dtype: str = "bf16"
    max_blocks_per_allocation_group: int = 64


class MemoryConfig(DeepSpeedConfigModel):
    mode: AllocationMode = AllocationMode.RESERVE
    size: int = Field(1_000_000_000, gt=0)  # RESERVE: bytes left free after the cache; ALLOCATE: number of blocks


class DSStateManagerConfig(DeepSpeedConfigModel):
    max_tracked_sequences: int = Field(2048, gt=0)
    max_ragged_batch_size: int = Field(768, gt=0)
    max_ragged_sequence_count: int = Field(512, gt=0)
    max_context: int = Field(8192, gt=0)
    memory_config: MemoryConfig = MemoryConfig()
    offload: bool = False
    # tokens per KV block; 0 = pick for the device (128 on CUDA: one block is one TMA page of the paged flash kernel; 16 on hosts).
    # The tensor-core attention kernel wants a divisor of 128.
    kv_block_size: int = Field(0, ge=0)

    @model_validator(mode="after")
    def max_ragged_sequence_count_validator(self):
        if self.max_ragged_sequence_count > self.max_tracked_sequences:
            raise ValueError("max_ragged_sequence_count must be less than max_tracked_sequences")
        if self.max_ragged_sequence_count > self.max_ragged_batch_size:
            raise ValueError("max_ragged_sequence_count must be less than max_ragged_batch_size")
        return self
