from .auto_tp import AutoTP, ReplaceWithTensorSlicing  # noqa: F401
from .layers import EmbeddingLayer, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, Normalize, OPTEmbedding, RMSNormalize  # noqa: F401
from .replace_module import GroupQuantizer, generic_injection, replace_transformer_layer, revert_transformer_layer  # noqa: F401
from .tp_shard import get_shard_size, get_shard_size_list, set_num_kv_heads  # noqa: F401
