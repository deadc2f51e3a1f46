"""(parity: inference/v2/model_implementations/mixtral/policy.py)"""
from ..family_policy import FamilyPolicy
from .container import MixtralNonTransformerContainer, MixtralTransformerContainer
from .model import MixtralInferenceModel


class MixtralPolicy(FamilyPolicy):
    model_types = ("mixtral", )
    model_class = MixtralInferenceModel
    transformer_container = MixtralTransformerContainer
    non_transformer_container = MixtralNonTransformerContainer
    transformer_prefix = "model.layers"
