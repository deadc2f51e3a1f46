"""(parity: inference/v2/model_implementations/opt/container.py)  LayerNorm with biases, learned positions (offset 2), ReLU MLP,
biases everywhere, tied unembedding."""
from typing import Optional

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, MLP1Parameter, MLP2Parameter, NormParameter,
                                 UnembedParameter, UnfusedQKVParameter)
from ..layer_container_base import LayerContainer


class OPTTransformerContainer(LayerContainer):
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
    norm2_w: NormParameter
    norm2_b: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.q_proj.bias": "qkv_b.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.k_proj.bias": "qkv_b.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.v_proj.bias": "qkv_b.v_params",
        "self_attn.out_proj.weight": "o_w.params",
        "self_attn.out_proj.bias": "o_b.params",
        "fc1.weight": "mlp1_w.params",
        "fc1.bias": "mlp1_b.params",
        "fc2.weight": "mlp2_w.params",
        "fc2.bias": "mlp2_b.params",
        "self_attn_layer_norm.weight": "norm1_w.params",
        "self_attn_layer_norm.bias": "norm1_b.params",
        "final_layer_norm.weight": "norm2_w.params",
        "final_layer_norm.bias": "norm2_b.params",
    }


class OPTNonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    pos_embed: EmbeddingParameter
    lm_head: UnembedParameter
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "model.decoder.embed_tokens.weight": "embed.params",
        "model.decoder.embed_positions.weight": "pos_embed.params",
        "model.decoder.final_layer_norm.weight": "final_norm_w.params",
        "model.decoder.final_layer_norm.bias": "final_norm_b.params",
        "lm_head.weight": "lm_head.params",
    }
