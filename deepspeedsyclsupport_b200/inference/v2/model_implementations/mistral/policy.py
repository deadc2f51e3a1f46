"""(parity: inference/v2/model_implementations/mistral/policy.py)"""
from ..family_policy import FamilyPolicy
from .container import MistralNonTransformerContainer, MistralTransformerContainer
from .model import MistralInferenceModel


class MistralPolicy(FamilyPolicy):
    model_types = ("mistral", )
    model_class = MistralInferenceModel
    transformer_container = MistralTransformerContainer
    non_transformer_container = MistralNonTransformerContainer
    transformer_prefix = "model.layers"
    tied_source = "model.embed_tokens.weight"
