from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ds_attention import DeepSpeedSelfAttention  # noqa: F401
from .ds_mlp import DeepSpeedMLP  # noqa: F401
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig  # noqa: F401,E402
