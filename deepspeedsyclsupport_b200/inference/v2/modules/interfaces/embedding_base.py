"""(parity: inference/v2/modules/interfaces/embedding_base.py:19-85)"""
from typing import Any, Dict, Optional, Type

import torch

from ...inference_parameter import InferenceParameter
from ..configs import DSEmbeddingsConfig
from ..ds_module import DSModuleBase
from ..module_registry import DSModuleRegistryBase


class DSEmbeddingBase(DSModuleBase):
    """Token (+ learned position, + token type) embedding lookup for the tokens of a ragged batch."""

    @staticmethod
    def config_class() -> Type[DSEmbeddingsConfig]:
        return DSEmbeddingsConfig

    def __init__(self, config: DSEmbeddingsConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)

    def transform_param(self, embed_param: torch.Tensor) -> InferenceParameter:
        raise NotImplementedError()

    @property
    def output(self) -> torch.Tensor:
        """Most recent output (kept for callers that want the buffer rather than the return value)."""
        raise NotImplementedError()

    def forward(self, ragged_batch, word_embeddings: torch.Tensor, position_embeddings: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None, token_type_embeddings: Optional[torch.Tensor] = None) -> torch.Tensor:
        raise NotImplementedError()


class DSEmbeddingRegistry(DSModuleRegistryBase):
    registry: Dict = {}

    @staticmethod
    def associated_class() -> Type[DSModuleBase]:
        return DSEmbeddingBase
