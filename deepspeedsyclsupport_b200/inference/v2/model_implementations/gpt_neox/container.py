"""GPT-NeoX / Pythia: per-head interleaved fused QKV, parallel residual with a separate LayerNorm for the MLP, partial rotary."""
from typing import Optional

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, MegatronQKVParameter, MLP1Parameter, MLP2Parameter,
                                 NormParameter, UnembedParameter)
from ..layer_container_base import LayerContainer


class GPTNeoXTransformerContainer(LayerContainer):
    qkv_w: MegatronQKVParameter
    qkv_b: Optional[MegatronQKVParameter]
    o_w: AttentionOutputParameter
    o_b: Optional[AttentionOutputParameter]
    mlp1_w: MLP1Parameter
    mlp1_b: Optional[MLP1Parameter]
    mlp2_w: MLP2Parameter
    mlp2_b: Optional[MLP2Parameter]
    norm1_w: NormParameter
    norm1_b: NormParameter
    norm2_w: NormParameter
    norm2_b: NormParameter

    PARAM_MAPPING = {
        "attention.query_key_value.weight": "qkv_w.params",
        "attention.query_key_value.bias": "qkv_b.params",
        "attention.dense.weight": "o_w.params",
        "attention.dense.bias": "o_b.params",
        "mlp.dense_h_to_4h.weight": "mlp1_w.params",
        "mlp.dense_h_to_4h.bias": "mlp1_b.params",
        "mlp.dense_4h_to_h.weight": "mlp2_w.params",
        "mlp.dense_4h_to_h.bias": "mlp2_b.params",
        "input_layernorm.weight": "norm1_w.params",
        "input_layernorm.bias": "norm1_b.params",
        "post_attention_layernorm.weight": "norm2_w.params",
        "post_attention_layernorm.bias": "norm2_b.params",
    }


class GPTNeoXNonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    lm_head: UnembedParameter
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "gpt_neox.embed_in.weight": "embed.params",
        "gpt_neox.final_layer_norm.weight": "final_norm_w.params",
        "gpt_neox.final_layer_norm.bias": "final_norm_b.params",
        "embed_out.weight": "lm_head.params",
    }
