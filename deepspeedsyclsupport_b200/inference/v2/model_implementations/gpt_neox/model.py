from ..inference_transformer_base import DSTransformerModelBase


class GPTNeoXInferenceModel(DSTransformerModelBase):
    pass
