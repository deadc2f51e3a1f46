"""(parity: inference/v2/model_implementations/llama_v2/llama_v2_model.py)  RMSNorm, rotary (rotate-half), gated SiLU MLP, GQA;
forward: RaggedTransformer.forward."""
from ..inference_transformer_base import DSTransformerModelBase


class Llama2InferenceModel(DSTransformerModelBase):
    pass
