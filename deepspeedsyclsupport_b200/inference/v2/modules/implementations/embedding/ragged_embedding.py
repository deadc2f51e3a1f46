"""(parity: inference/v2/modules/implementations/embedding/ragged_embedding.py:20-77)"""
from typing import Any, Dict, Optional

import torch

from .... import kernels as K
from ....inference_parameter import InferenceParameter
from ....inference_utils import DtypeEnum
from ...configs import DSEmbeddingsConfig
from ...interfaces import DSEmbeddingBase, DSEmbeddingRegistry


@DSEmbeddingRegistry.register_module
class DSRaggedEmbedding(DSEmbeddingBase):

    @staticmethod
    def name() -> str:
        return "ragged_embedding"

    @staticmethod
    def supports_config(config: DSEmbeddingsConfig) -> bool:
        if DtypeEnum(config.residual_dtype) not in (DtypeEnum.fp16, DtypeEnum.bf16, DtypeEnum.fp32):
            return False
        return not config.use_token_type and config.output_normalization is None

    def __init__(self, config: DSEmbeddingsConfig, implementation_config: Dict[str, Any] = None) -> None:
        super().__init__(config, implementation_config)
        self._output = None

    def transform_param(self, embed_param: torch.Tensor) -> InferenceParameter:
        return InferenceParameter.initialize(embed_param.to(DtypeEnum(self._config.residual_dtype).value).contiguous())

    @property
    def output(self) -> torch.Tensor:
        return self._output

    def forward(self, ragged_batch, word_embeddings: torch.Tensor, position_embeddings: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None, token_type_embeddings: Optional[torch.Tensor] = None) -> torch.Tensor:
        assert token_type_ids is None and token_type_embeddings is None, "token-type embeddings are not supported"
        self._output = K.ragged_embed(ragged_batch.input_ids(), word_embeddings, position_embeddings,
                                      ragged_batch.tok_pos() if position_embeddings is not None else None,
                                      self._config.positional_offset)
        return self._output
