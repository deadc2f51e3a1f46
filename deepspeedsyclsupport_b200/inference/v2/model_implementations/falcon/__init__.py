from .container import FalconNonTransformerContainer, FalconTransformerContainer
from .model import FalconInferenceModel
from .policy import FalconPolicy
