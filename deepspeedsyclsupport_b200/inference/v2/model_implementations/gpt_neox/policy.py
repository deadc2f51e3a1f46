from ..family_policy import FamilyPolicy
from .container import GPTNeoXNonTransformerContainer, GPTNeoXTransformerContainer
from .model import GPTNeoXInferenceModel


class GPTNeoXPolicy(FamilyPolicy):
    model_types = ("gpt_neox", )
    model_class = GPTNeoXInferenceModel
    transformer_container = GPTNeoXTransformerContainer
    non_transformer_container = GPTNeoXNonTransformerContainer
    transformer_prefix = "gpt_neox.layers"
