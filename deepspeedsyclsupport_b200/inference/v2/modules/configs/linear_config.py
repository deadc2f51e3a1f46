"""(parity: inference/v2/modules/configs/linear_config.py:12-43)"""
from ...inference_utils import ActivationType, DtypeEnum
from ..ds_module import DSModuleConfig


class DSLinearConfig(DSModuleConfig):
    in_channels: int
    out_channels: int  # with a gated activation the weight has 2 * out_channels rows
    activation: ActivationType = ActivationType.IDENTITY
    input_dtype: DtypeEnum = DtypeEnum.fp16
    output_dtype: DtypeEnum = DtypeEnum.fp16
