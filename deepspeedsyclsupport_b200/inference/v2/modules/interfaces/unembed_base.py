"""(parity: inference/v2/modules/interfaces/unembed_base.py:17-61)"""
from typing import Any, Dict, Optional, Type

import torch

from ..configs import DSUnembedConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSUnembedBase(DSModuleBase):
    """Gather the last token of every sequence, (optionally) final norm, project onto the (local shard of the) vocabulary."""

    @staticmethod
    def config_class() -> Type[DSUnembedConfig]:
        return DSUnembedConfig

    def __init__(self, config: DSUnembedConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    def forward(self, hidden_states: torch.Tensor, vocab_embedding: torch.Tensor, ragged_metadata, bias: Optional[torch.Tensor] = None,
                gamma: Optional[torch.Tensor] = None, beta: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError()


class DSUnembedRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSUnembedBase
