from ..inference_transformer_base import DSTransformerModelBase


class GPTJInferenceModel(DSTransformerModelBase):
    pass
