from .container import MistralNonTransformerContainer, MistralTransformerContainer
from .model import MistralInferenceModel
from .policy import MistralPolicy
