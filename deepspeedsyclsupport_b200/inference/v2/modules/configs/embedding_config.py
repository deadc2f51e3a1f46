"""(parity: inference/v2/modules/configs/embedding_config.py:14-70)"""
from typing import Optional

from ...inference_utils import DtypeEnum, NormTypeEnum
from ..ds_module import DSModuleConfig


class DSEmbeddingsConfig(DSModuleConfig):
    residual_dtype: DtypeEnum = DtypeEnum.fp16
    embedding_dim: int
    positional_embedding: bool = False
    positional_offset: int = 0  # OPT stores position p at row p + 2
    use_token_type: bool = False
    output_normalization: Optional[NormTypeEnum] = None
