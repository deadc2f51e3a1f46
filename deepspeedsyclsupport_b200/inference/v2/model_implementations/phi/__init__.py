from .container import PhiNonTransformerContainer, PhiTransformerContainer
from .model import PhiInferenceModel
from .policy import PhiPolicy
