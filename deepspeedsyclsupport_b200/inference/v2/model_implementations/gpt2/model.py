from ..inference_transformer_base import DSTransformerModelBase


class GPT2InferenceModel(DSTransformerModelBase):
    pass
