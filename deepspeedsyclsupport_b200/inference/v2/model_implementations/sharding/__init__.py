from .attn import get_local_heads
from .attn_out import attn_out_in_features, shard_attn_out_param
from .embedding import shard_embedding_param, sharded_embedding_dim
from .mlp import shard_mlp_1_param, shard_mlp_2_param, sharded_intermediate_dim
from .qkv import qkv_out_features, shard_qkv_param
from .types import DEFAULT_SHARD_GRANULARITY, ShardingType
from .unembed import shard_unembed_param, sharded_unembed_dim
from .utils import get_shard_endpoints, shard_param
