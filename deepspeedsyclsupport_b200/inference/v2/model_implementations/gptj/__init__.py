from .container import GPTJNonTransformerContainer, GPTJTransformerContainer
from .model import GPTJInferenceModel
from .policy import GPTJPolicy
