"""Per-sequence tracking state (parity: inference/v2/ragged/sequence_descriptor.py:59-276)."""
import torch


class DSSequenceDescriptor:

    def __init__(self, tracking_id, max_blocks, block_size):
        self.tracking_id = tracking_id
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
