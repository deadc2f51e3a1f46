from .attn_parameters import (AttentionOutputParameter, FusedQKVParameter, GQAMegatronQKVParameter, MegatronQKVParameter, UnfusedQKVParameter,
                              gqa_megatron_qkv_reshape, megatron_qkv_reshape)
from .embedding_parameters import EmbeddingParameter
from .invfreq_parameters import InvFreqParameter
from .mlp_parameters import GatedMLPParameter, MLP1Parameter, MLP2Parameter
from .moe_parameters import MoEGatingWeightParameter, UnfusedMoEGatedMLPParameter, UnfusedMoEMLP1Parameter, UnfusedMoEMLP2Parameter
from .norm_parameters import NormParameter
from .unembed_parameters import UnembedParameter
