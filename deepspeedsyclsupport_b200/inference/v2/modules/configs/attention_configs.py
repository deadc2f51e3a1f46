"""(parity: inference/v2/modules/configs/attention_configs.py:16-110)"""
from enum import Enum
from typing import Dict, Optional

from .....runtime.config_utils import DeepSpeedConfigModel
from ...inference_utils import DtypeEnum
from ..ds_module import DSModuleConfig


class PositionalEmbeddingType(Enum):
    none = "none"
    rotate_half = "rotate_half"  # rotary, (x[:d/2], x[d/2:]) pairing (llama / neox style)
    rotate_every_other = "rotate_every_other"  # rotary, (x[2i], x[2i+1]) pairing (gpt-j style)
    alibi = "alibi"


class RotateHalfConfig(DeepSpeedConfigModel):
    use_trained_freqs: bool = False  # take `inv_freqs` from the caller instead of synthesising theta_base ** (-2i/d)
    theta_base: float = 10_000.0
    rotate_dim: Optional[int] = None  # number of rotated channels per head; None = all (no hidden doubling)


class MaskingType(Enum):
    none = "none"
    causal = "causal"
    local = "local"
    symmetric = "symmetric"
    asymmetric = "asymmetric"


class DSSelfAttentionConfig(DSModuleConfig):
    n_heads_q: int
    n_heads_kv: int
    head_size: int
    max_sequences: int
    scale_factor: float = 1.0
    input_dtype: DtypeEnum = DtypeEnum.fp16
    output_dtype: DtypeEnum = DtypeEnum.fp16
    masking_type: MaskingType = MaskingType.causal
    masking_args: Dict = {}
    positional_embedding_type: PositionalEmbeddingType = PositionalEmbeddingType.none
    positional_embedding_config: Optional[RotateHalfConfig] = None
