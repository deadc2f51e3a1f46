"""(parity: inference/v2/model_implementations/llama_v2/llama_v2_policy.py)"""
from ..family_policy import FamilyPolicy
from .container import Llama2NonTransformerContainer, Llama2TransformerContainer
from .model import Llama2InferenceModel


class Llama2Policy(FamilyPolicy):
    model_types = ("llama", "internlm")
    model_class = Llama2InferenceModel
    transformer_container = Llama2TransformerContainer
    non_transformer_container = Llama2NonTransformerContainer
    transformer_prefix = "model.layers"
    tied_source = "model.embed_tokens.weight"
