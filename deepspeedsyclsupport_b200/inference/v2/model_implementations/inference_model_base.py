"""What the FastGen engine needs from a model (parity: inference/v2/model_implementations/inference_model_base.py:27-257).
`RaggedTransformer` (ragged_transformer.py) is the one concrete transformer implementation; the per-family classes under
model_implementations/<family>/model.py bind it to that family's config translation and containers."""
from abc import ABC, abstractmethod
from typing import Iterable, Optional, Tuple

import torch

MAX_INLINE_SEQ = 1


class DSInferenceModelBase(ABC):

    @property
    @abstractmethod
    def tp_rank(self) -> int:
        ...

    @property
    @abstractmethod
    def tp_size(self) -> int:
        ...

    @abstractmethod
    def named_containers(self) -> Iterable[Tuple[str, object]]:
        """(name, LayerContainer) of every container holding parameters of this rank."""

    @abstractmethod
    def forward(self, batch, state_manager) -> torch.Tensor:
        """Logits of the last token of every sequence in the ragged batch."""

    # ---- KV-cache requirements, used by the state manager to size the cache and by the scheduler to admit work
    def kv_cache_config(self):
        from ..ragged.manager_configs import KVCacheConfig
        s = self.spec
        name = {torch.bfloat16: "bf16", torch.float16: "fp16", torch.float32: "fp32"}[self.dtype]
        return KVCacheConfig(block_size=128, cache_shape=(s.num_layers, self.hkv, s.head_dim), cache_dtype=name,
                             max_blocks_per_allocation_group=-(-s.max_positions // 128))

    def get_kv_requirements(self, sequence, max_new_tokens: int, max_new_blocks: int, block_size: int = 128) -> Tuple[int, int]:
        """How many of `max_new_tokens` fit if at most `max_new_blocks` more blocks may be given to `sequence`
        -> (schedulable tokens, blocks needed for them)."""
        seen = sequence.seen_tokens
        held = sequence.cur_allocated_blocks
        free_in_held = held * block_size - seen
        tokens = min(max_new_tokens, free_in_held + max_new_blocks * block_size)
        blocks = max(0, -(-(seen + tokens) // block_size) - held)
        return tokens, blocks

    def maybe_allocate_kv(self, sequence, n_new_tokens: int, state_manager, block_size: int = 128) -> None:
        _, blocks = self.get_kv_requirements(sequence, n_new_tokens, 1 << 30, block_size)
        if blocks:
            sequence.extend_kv_cache(state_manager.allocate_blocks(blocks))

    def get_remaining_block_capacity(self, sequence, block_size: int = 128) -> int:
        return sequence.cur_allocated_blocks * block_size - sequence.seen_tokens

    def maybe_free_kv(self, sequence) -> None:
        """Sliding-window models release blocks that fell out of the window; full-attention models keep everything."""
