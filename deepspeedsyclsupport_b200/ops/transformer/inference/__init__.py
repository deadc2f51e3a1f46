from .config import DeepSpeedInferenceConfig  # noqa: F401
from .ds_attention import DeepSpeedSelfAttention  # noqa: F401
from .ds_mlp import DeepSpeedMLP  # noqa: F401
from .moe_inference import DeepSpeedMoEInference, DeepSpeedMoEInferenceConfig  # noqa: F401,E402
from .bias_add import nhwc_bias_add  # noqa: F401,E402
from .diffusers_2d_transformer import Diffusers2DTransformerConfig  # noqa: F401,E402
from .diffusers_attention import DeepSpeedDiffusersAttention  # noqa: F401,E402
from .diffusers_transformer_block import DeepSpeedDiffusersTransformerBlock  # noqa: F401,E402
