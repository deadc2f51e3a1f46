"""Configuration of the pre- / post-norm modules (parity: inference/v2/modules/configs/norm_config.py:12-32)."""
from pydantic import field_validator

from ...inference_utils import DtypeEnum, NormTypeEnum
from ..ds_module import DSModuleConfig


class DSNormConfig(DSModuleConfig):
    """One config serves DSPreNormBase and DSPostNormBase.  `residual_dtype` is the dtype of the residual stream the norm adds into,
    `input_dtype` that of the incoming hidden states, `output_dtype` that of the normalised result."""
    type: NormTypeEnum  # layer_norm | rms_norm
    channels: int  # width of the normalised (last) dimension
    eps: float = 1e-5
    residual_dtype: DtypeEnum = DtypeEnum.fp16
    input_dtype: DtypeEnum = DtypeEnum.fp16
    output_dtype: DtypeEnum = DtypeEnum.fp16

    @field_validator("channels")
    @classmethod
    def _positive_width(cls, v):
        if v <= 0:
            raise ValueError(f"channels must be positive, got {v}")
        return v

    @field_validator("eps")
    @classmethod
    def _positive_eps(cls, v):
        if not v > 0:
            raise ValueError(f"eps must be positive, got {v}")
        return v

    @property
    def is_rms(self) -> bool:
        return NormTypeEnum(self.type) == NormTypeEnum.RMSNorm
