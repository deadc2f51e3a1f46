from .attention_configs import DSSelfAttentionConfig, MaskingType, PositionalEmbeddingType, RotateHalfConfig
from .embedding_config import DSEmbeddingsConfig
from .linear_config import DSLinearConfig
from .moe_config import DSMoEConfig
from .norm_config import DSNormConfig
from .unembed_config import DSUnembedConfig
