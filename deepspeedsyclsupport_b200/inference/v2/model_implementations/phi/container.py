"""(parity: inference/v2/model_implementations/phi/containers.py)  Parallel block with one LayerNorm, partial rotary, GELU MLP,
biases on every projection including the unembedding."""
from typing import Optional

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, MLP1Parameter, MLP2Parameter, NormParameter,
                                 UnembedParameter, UnfusedQKVParameter)
from ..layer_container_base import LayerContainer


class PhiTransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    qkv_b: UnfusedQKVParameter
    o_w: AttentionOutputParameter
    o_b: AttentionOutputParameter
    mlp1_w: MLP1Parameter
    mlp1_b: MLP1Parameter
    mlp2_w: MLP2Parameter
    mlp2_b: MLP2Parameter
    norm1_w: NormParameter
    norm1_b: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.q_proj.bias": "qkv_b.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.k_proj.bias": "qkv_b.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.v_proj.bias": "qkv_b.v_params",
        "self_attn.dense.weight": "o_w.params",
        "self_attn.dense.bias": "o_b.params",
        "mlp.fc1.weight": "mlp1_w.params",
        "mlp.fc1.bias": "mlp1_b.params",
        "mlp.fc2.weight": "mlp2_w.params",
        "mlp.fc2.bias": "mlp2_b.params",
        "input_layernorm.weight": "norm1_w.params",
        "input_layernorm.bias": "norm1_b.params",
    }


class PhiNonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    lm_head: UnembedParameter
    lm_head_b: Optional[UnembedParameter]
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "model.embed_tokens.weight": "embed.params",
        "model.final_layernorm.weight": "final_norm_w.params",
        "model.final_layernorm.bias": "final_norm_b.params",
        "lm_head.weight": "lm_head.params",
        "lm_head.bias": "lm_head_b.params",
    }
