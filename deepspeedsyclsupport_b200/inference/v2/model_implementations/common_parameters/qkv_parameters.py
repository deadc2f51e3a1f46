"""(parity: inference/v2/model_implementations/common_parameters/qkv_parameters.py)  The QKV parameter flavours live in
attn_parameters.py next to the attention-output parameter."""
from .attn_parameters import (FusedQKVParameter, GQAMegatronQKVParameter, MegatronQKVParameter, UnfusedQKVParameter,  # noqa: F401
                              gqa_megatron_qkv_reshape, megatron_qkv_reshape)
