"""(parity: inference/v2/modules/configs/unembed_config.py:12-39)"""
from typing import Optional

from ...inference_utils import DtypeEnum, NormTypeEnum
from ..ds_module import DSModuleConfig


class DSUnembedConfig(DSModuleConfig):
    dtype: DtypeEnum = DtypeEnum.fp16
    norm_type: Optional[NormTypeEnum] = None
    model_dim: int
    max_sequences: int
    vocab_size: int
    eps: float = 1e-5
