from .container import GPT2NonTransformerContainer, GPT2TransformerContainer
from .model import GPT2InferenceModel
from .policy import GPT2Policy
