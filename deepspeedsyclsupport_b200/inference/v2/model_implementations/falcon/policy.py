"""(parity: inference/v2/model_implementations/falcon/policy.py)"""
from ..family_policy import FamilyPolicy
from .container import FalconNonTransformerContainer, FalconTransformerContainer
from .model import FalconInferenceModel


class FalconPolicy(FamilyPolicy):
    model_types = ("falcon", )
    model_class = FalconInferenceModel
    transformer_container = FalconTransformerContainer
    non_transformer_container = FalconNonTransformerContainer
    transformer_prefix = "transformer.h"
    tied_source = "transformer.word_embeddings.weight"
