"""(parity: inference/v2/model_implementations/mistral/model.py)  Llama architecture with grouped-query attention; the
sliding window is longer than the contexts the engine admits (max_context), so attention is plain causal."""
from ..inference_transformer_base import DSTransformerModelBase


class MistralInferenceModel(DSTransformerModelBase):
    pass
