"""Configuration of the self-attention modules (parity: inference/v2/modules/configs/attention_configs.py:16-110)."""
from enum import Enum
from typing import Dict, Optional

from pydantic import model_validator

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

    @model_validator(mode="after")
    def _consistent_heads(self):
        if self.n_heads_q <= 0 or self.n_heads_kv <= 0 or self.head_size <= 0:
            raise ValueError("head counts and head size must be positive")
        if self.n_heads_q % self.n_heads_kv != 0:
            raise ValueError(f"n_heads_q ({self.n_heads_q}) must be a multiple of n_heads_kv ({self.n_heads_kv})")
        rc = self.positional_embedding_config
        if rc is not None and rc.rotate_dim is not None and not (0 < rc.rotate_dim <= self.head_size and rc.rotate_dim % 2 == 0):
            raise ValueError(f"rotate_dim must be even and at most head_size ({self.head_size}), got {rc.rotate_dim}")
        return self

    @property
    def group_size(self) -> int:
        """query heads served by one key/value head"""
        return self.n_heads_q // self.n_heads_kv
