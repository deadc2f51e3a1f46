from .container import MixtralNonTransformerContainer, MixtralTransformerContainer
from .model import MixtralInferenceModel
from .policy import MixtralPolicy
