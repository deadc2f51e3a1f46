"""Mistral stores its tensors under the Llama names (parity: inference/v2/model_implementations/mistral/container.py)."""
from ..llama_v2.container import Llama2NonTransformerContainer, Llama2TransformerContainer


class MistralTransformerContainer(Llama2TransformerContainer):
    pass


class MistralNonTransformerContainer(Llama2NonTransformerContainer):
    pass
