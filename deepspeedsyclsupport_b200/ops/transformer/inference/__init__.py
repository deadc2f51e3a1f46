from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ds_attention import DeepSpeedSelfAttention  # noqa: F401
from .ds_mlp import DeepSpeedMLP  # noqa: F401
