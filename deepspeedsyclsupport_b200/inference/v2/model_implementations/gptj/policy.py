from ..family_policy import FamilyPolicy
from .container import GPTJNonTransformerContainer, GPTJTransformerContainer
from .model import GPTJInferenceModel


class GPTJPolicy(FamilyPolicy):
    model_types = ("gptj", )
    model_class = GPTJInferenceModel
    transformer_container = GPTJTransformerContainer
    non_transformer_container = GPTJNonTransformerContainer
    transformer_prefix = "transformer.h"
