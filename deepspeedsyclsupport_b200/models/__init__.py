from .llama import LlamaConfig, LlamaDecoderLayer, LlamaForCausalLM, LlamaModel, llama_config  # noqa: F401
