from .container import GPTNeoXNonTransformerContainer, GPTNeoXTransformerContainer
from .model import GPTNeoXInferenceModel
from .policy import GPTNeoXPolicy
