from .transformer import DeepSpeedTransformerConfig, DeepSpeedTransformerLayer, TransformerConfig
from .inference.config import DeepSpeedInferenceConfig  # noqa: F401,E402
from .inference.moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig  # noqa: F401,E402


def __getattr__(name):
    # resolved lazily: the block implementation imports this package's inference ops
    if name == "DeepSpeedTransformerInference":
        from ...model_implementations.transformers.ds_transformer import DeepSpeedTransformerInference
        return DeepSpeedTransformerInference
    raise AttributeError(name)
