"""(parity: inference/v2/model_implementations/falcon/model.py)"""
from ..inference_transformer_base import DSTransformerModelBase


class FalconInferenceModel(DSTransformerModelBase):
    pass
