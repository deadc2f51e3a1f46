from .gated_mlp import HybridGatedMLPContainer
from .hybrid_engine import HybridEngineContainer
from .hybrid_megatron import HybridMegatronContainer
from .megatron import MegatronContainer
from .meta_tensor import MetaTensorContainer
from .split_qkv import HybridSplitQKVContainer
