from .auto_tp import AutoTP, ReplaceWithTensorSlicing  # noqa: F401
from .layers import EmbeddingLayer, LinearAllreduce, LinearLayer, LmHeadLinearAllreduce, Normalize, OPTEmbedding, RMSNormalize  # noqa: F401
from .replace_module import GroupQuantizer, generic_injection, replace_transformer_layer, revert_transformer_layer  # noqa: F401
from .tp_shard import get_shard_size, get_shard_size_list, set_num_kv_heads  # noqa: F401
from .inject import module_inject  # noqa: F401
from .load_checkpoint import load_model_with_checkpoint  # noqa: F401
from .module_quantize import quantize_transformer_layer  # noqa: F401
from .policy import DSPolicy, TransformerPolicy  # noqa: F401
from .replace_module import replace_with_kernel_policies  # noqa: F401
from .replace_policy import generic_policies, replace_policies  # noqa: F401
from .utils import policy_to_ds_container  # noqa: F401
from .containers import HFBertLayerPolicy  # noqa: F401,E402
