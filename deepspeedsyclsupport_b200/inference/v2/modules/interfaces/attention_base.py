"""(parity: inference/v2/modules/interfaces/attention_base.py:17-97)"""
from typing import Any, Dict, Optional, Type

import torch

from ..configs import DSSelfAttentionConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSSelfAttentionBase(DSModuleBase):
    """Self-attention over a ragged batch against the blocked KV cache: consumes fused [tokens, (hq + 2 hkv) * d] projections,
    appends K/V (after the positional rotation, if any) to the cache and returns [tokens, hq * d]."""

    @staticmethod
    def config_class() -> Type[DSSelfAttentionConfig]:
        return DSSelfAttentionConfig

    def __init__(self, config: DSSelfAttentionConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    @property
    def kv_block_size(self) -> int:
        raise NotImplementedError()

    @property
    def q_block_size(self) -> int:
        raise NotImplementedError()

    def build_atoms(self, ragged_batch) -> None:
        """Per-batch work-list construction, shared by every layer (optional)."""

    def forward(self, q_k_v: torch.Tensor, kv_cache: torch.Tensor, batch, attention_mask: Optional[torch.Tensor] = None,
                attention_bias: Optional[torch.Tensor] = None, inv_freqs: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError()


class DSSelfAttentionRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSSelfAttentionBase
