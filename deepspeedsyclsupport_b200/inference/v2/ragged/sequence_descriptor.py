"""Per-sequence tracking state (parity: inference/v2/ragged/sequence_descriptor.py:59-276)."""
import torch


class DSSequenceDescriptor:

    def __init__(self, tracking_id, max_blocks, block_size):
        self._tracking_id = tracking_id
        self._seen_tokens = 0
        self._in_flight_tokens = 0
        self._block_size = block_size
        self._blocks = torch.zeros(max_blocks, dtype=torch.int32)
        self._num_blocks = 0

    @property
    def seen_tokens(self):
        return self._seen_tokens

    @property
    def in_flight_tokens(self):
        return self._in_flight_tokens

    @property
    def cur_allocated_blocks(self):
        return self._num_blocks

    @property
    def max_blocks(self):
        return self._blocks.numel()

    def kv_blocks_needed(self, new_tokens):
        total = self._seen_tokens + new_tokens
        return max(0, -(-total // self._block_size) - self._num_blocks)

    def extend_kv_cache(self, new_ids):
        n = len(new_ids)
        assert self._num_blocks + n <= self._blocks.numel(), "sequence exceeds max_context"
        self._blocks[self._num_blocks:self._num_blocks + n] = torch.as_tensor(new_ids, dtype=torch.int32)
        self._num_blocks += n

    def kv_block_ids(self):
        return self._blocks[:self._num_blocks]

    @property
    def all_block_ids(self):
        return self._blocks[:self._num_blocks]

    def pre_forward(self, num_tokens):
        self._in_flight_tokens = num_tokens

    def post_forward(self):
        self._seen_tokens += self._in_flight_tokens
        self._in_flight_tokens = 0

    @property
    def tracking_id(self):
        return self._tracking_id

    @property
    def max_context(self):
        """Longest context this descriptor can describe (its block-table row is full beyond that)."""
        return self._blocks.numel() * self._block_size

    def kv_cache_ids(self, cache_group=0, on_device=False):
        """Block ids of the sequence (one cache group).  on_device: the row as a device tensor for kernels."""
        ids = self._blocks[:self._num_blocks]
        if on_device:
            from ....accelerator import get_accelerator
            return ids.to(get_accelerator().current_device_name())
        return ids

    def kv_blocks_ptr(self, cache_group=0):
        """Address of the block-id row (host memory); kernels here take the batch block table instead, kept for API parity."""
        return self._blocks.data_ptr()

    def free_kv_cache(self, free_ids, cache_group=0):
        """Drop the given block ids from this sequence's table (sliding-window models release blocks that fell out of the
        window).  Ids must currently belong to the sequence; order of the remaining blocks is preserved."""
        drop = set(int(i) for i in (free_ids.tolist() if torch.is_tensor(free_ids) else free_ids))
        kept = [int(b) for b in self._blocks[:self._num_blocks].tolist() if b not in drop]
        assert len(kept) == self._num_blocks - len(drop), "free_kv_cache: ids not owned by the sequence"
        self._blocks[:len(kept)] = torch.as_tensor(kept, dtype=torch.int32) if kept else self._blocks[:0]
        self._num_blocks = len(kept)
