"""(parity: inference/v2/model_implementations/opt/policy.py)"""
from ..family_policy import FamilyPolicy
from .container import OPTNonTransformerContainer, OPTTransformerContainer
from .model import OPTInferenceModel


class OPTPolicy(FamilyPolicy):
    model_types = ("opt", )
    model_class = OPTInferenceModel
    transformer_container = OPTTransformerContainer
    non_transformer_container = OPTNonTransformerContainer
    transformer_prefix = "model.decoder.layers"
    tied_source = "model.decoder.embed_tokens.weight"

    def rename(self, name):  # state dicts taken from OPTModel lack the "model." prefix
        return "model." + name if name.startswith("decoder.") else name
