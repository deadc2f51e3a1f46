"""(parity: inference/v2/model_implementations/opt/model.py)"""
from ..inference_transformer_base import DSTransformerModelBase


class OPTInferenceModel(DSTransformerModelBase):
    pass
