"""(parity: inference/v2/model_implementations/phi/model.py)"""
from ..inference_transformer_base import DSTransformerModelBase


class PhiInferenceModel(DSTransformerModelBase):
    pass
