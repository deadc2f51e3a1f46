"""DSStateManager: tracked sequences + KV cache (parity: inference/v2/ragged/ragged_manager.py:19-206)."""
from typing import Dict, Optional

import torch

from ....utils.logging import logger
from .kv_cache import BlockedKVCache
from .manager_configs import DSStateManagerConfig, KVCacheConfig
from .sequence_descriptor import DSSequenceDescriptor


class DSStateManager:

    def __init__(self, config: DSStateManagerConfig, kv_config: KVCacheConfig, base_mp_group=None):
        self._config = config
        self._kv_config = kv_config
        self._seqs: Dict[int, DSSequenceDescriptor] = {}
        self._kv_cache = BlockedKVCache(kv_config, config.memory_config, mp_group=base_mp_group, offload=config.offload)
        self.max_blocks_per_seq = -(-config.max_context // kv_config.block_size)

    def get_cache(self, layer):
        return self._kv_cache.get_cache(layer)

    def query(self, uid: Optional[int] = None):
        if uid is None:
            return self._kv_cache.free_blocks
        return self._seqs.get(uid, None)

    def get_sequence(self, uid):
        return self._seqs.get(uid, None)

    def get_or_create_sequence(self, uid):
        seq = self.get_sequence(uid)
        if seq is not None:
            return seq
        return self._create_sequence(uid)

    def _create_sequence(self, uid):
        if uid in self._seqs:
            raise ValueError(f"Sequence {uid} already exists.")
        if len(self._seqs) >= self._config.max_tracked_sequences:
            raise RuntimeError(f"Too many tracked sequences (> {self._config.max_tracked_sequences})")
        self._seqs[uid] = DSSequenceDescriptor(uid, self.max_blocks_per_seq, self._kv_config.block_size)
        return self._seqs[uid]

    def flush_sequence(self, uid):
        if uid not in self._seqs:
            logger.warning(f"Attempting to flush sequence {uid} which does not exist.")
            return
        seq = self._seqs.pop(uid)
        if seq.cur_allocated_blocks > 0:
            self._kv_cache.free(seq.all_block_ids.clone())

    def allocate_blocks(self, n_blocks):
        return self._kv_cache.reserve(n_blocks)

    @property
    def tracked_sequences(self):
        return self._seqs

    @property
    def n_tracked_sequences(self):
        return len(self._seqs)

    @property
    def kv_block_size(self):
        return self._kv_config.block_size

    @property
    def free_blocks(self):
        return self._kv_cache.free_blocks

    @property
    def n_kv_cache_groups(self):
        return self._kv_cache.num_caches
