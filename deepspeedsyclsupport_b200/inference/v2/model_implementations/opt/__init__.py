from .container import OPTNonTransformerContainer, OPTTransformerContainer
from .model import OPTInferenceModel
from .policy import OPTPolicy
