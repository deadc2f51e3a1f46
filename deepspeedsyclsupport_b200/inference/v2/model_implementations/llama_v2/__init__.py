from .container import Llama2NonTransformerContainer, Llama2TransformerContainer
from .model import Llama2InferenceModel
from .policy import Llama2Policy
