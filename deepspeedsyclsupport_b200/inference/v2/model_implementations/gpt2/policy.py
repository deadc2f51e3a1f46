from ..family_policy import FamilyPolicy
from .container import GPT2NonTransformerContainer, GPT2TransformerContainer
from .model import GPT2InferenceModel


class GPT2Policy(FamilyPolicy):
    model_types = ("gpt2", )
    model_class = GPT2InferenceModel
    transformer_container = GPT2TransformerContainer
    non_transformer_container = GPT2NonTransformerContainer
    transformer_prefix = "transformer.h"
    tied_source = "transformer.wte.weight"

    def rename(self, name):  # state dicts taken from GPT2Model lack the "transformer." prefix
        return name if name.startswith(("transformer.", "lm_head.")) else "transformer." + name
