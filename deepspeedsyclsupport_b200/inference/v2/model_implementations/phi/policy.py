"""(parity: inference/v2/model_implementations/phi/policy.py)"""
from ..family_policy import FamilyPolicy
from .container import PhiNonTransformerContainer, PhiTransformerContainer
from .model import PhiInferenceModel


class PhiPolicy(FamilyPolicy):
    model_types = ("phi", )
    model_class = PhiInferenceModel
    transformer_container = PhiTransformerContainer
    non_transformer_container = PhiNonTransformerContainer
    transformer_prefix = "model.layers"
