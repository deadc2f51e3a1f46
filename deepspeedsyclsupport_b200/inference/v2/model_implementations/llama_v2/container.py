"""Llama-2 tensors (HF naming) -> parameters (parity: inference/v2/model_implementations/llama_v2/llama_v2_containers.py)."""
from typing import Optional

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, GatedMLPParameter, MLP2Parameter, NormParameter,
                                 UnembedParameter, UnfusedQKVParameter)
from ..layer_container_base import LayerContainer


class Llama2TransformerContainer(LayerContainer):
    qkv_w: UnfusedQKVParameter
    qkv_b: Optional[UnfusedQKVParameter]  # internlm / attention_bias checkpoints
    o_w: AttentionOutputParameter
    o_b: Optional[AttentionOutputParameter]
    mlp1_w: GatedMLPParameter
    mlp2_w: MLP2Parameter
    norm1_w: NormParameter
    norm2_w: NormParameter

    PARAM_MAPPING = {
        "self_attn.q_proj.weight": "qkv_w.q_params",
        "self_attn.k_proj.weight": "qkv_w.k_params",
        "self_attn.v_proj.weight": "qkv_w.v_params",
        "self_attn.q_proj.bias": "qkv_b.q_params",
        "self_attn.k_proj.bias": "qkv_b.k_params",
        "self_attn.v_proj.bias": "qkv_b.v_params",
        "self_attn.o_proj.weight": "o_w.params",
        "self_attn.o_proj.bias": "o_b.params",
        "mlp.gate_proj.weight": "mlp1_w.gate_params",
        "mlp.up_proj.weight": "mlp1_w.up_params",
        "mlp.down_proj.weight": "mlp2_w.params",
        "input_layernorm.weight": "norm1_w.params",
        "post_attention_layernorm.weight": "norm2_w.params",
    }


class Llama2NonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    lm_head: UnembedParameter
    final_norm_w: NormParameter

    PARAM_MAPPING = {
        "model.embed_tokens.weight": "embed.params",
        "model.norm.weight": "final_norm_w.params",
        "lm_head.weight": "lm_head.params",
    }
