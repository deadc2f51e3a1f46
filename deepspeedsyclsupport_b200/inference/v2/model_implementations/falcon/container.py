"""(parity: inference/v2/model_implementations/falcon/container.py)  The fused query_key_value tensor is grouped per KV head:
[group's query heads, k, v] -- one group for multi-query (7b), several for 40b/180b, group size 1 when every head has its own
KV (falcon-rw).  Parallel attention/MLP block; the new decoder architecture adds a separate LayerNorm for the MLP."""
from typing import Optional

from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, GQAMegatronQKVParameter, MLP1Parameter, MLP2Parameter,
                                 NormParameter, UnembedParameter)
from ..layer_container_base import LayerContainer


class FalconTransformerContainer(LayerContainer):
    qkv_w: GQAMegatronQKVParameter
    qkv_b: Optional[GQAMegatronQKVParameter]
    o_w: AttentionOutputParameter
    o_b: Optional[AttentionOutputParameter]
    mlp1_w: MLP1Parameter
    mlp1_b: Optional[MLP1Parameter]
    mlp2_w: MLP2Parameter
    mlp2_b: Optional[MLP2Parameter]
    norm1_w: NormParameter
    norm1_b: NormParameter
    norm2_w: Optional[NormParameter]
    norm2_b: Optional[NormParameter]

    PARAM_MAPPING = {
        "self_attention.query_key_value.weight": "qkv_w.params",
        "self_attention.query_key_value.bias": "qkv_b.params",
        "self_attention.dense.weight": "o_w.params",
        "self_attention.dense.bias": "o_b.params",
        "mlp.dense_h_to_4h.weight": "mlp1_w.params",
        "mlp.dense_h_to_4h.bias": "mlp1_b.params",
        "mlp.dense_4h_to_h.weight": "mlp2_w.params",
        "mlp.dense_4h_to_h.bias": "mlp2_b.params",
        "input_layernorm.weight": "norm1_w.params",
        "input_layernorm.bias": "norm1_b.params",
        "ln_attn.weight": "norm1_w.params",
        "ln_attn.bias": "norm1_b.params",
        "ln_mlp.weight": "norm2_w.params",
        "ln_mlp.bias": "norm2_b.params",
        "post_attention_layernorm.weight": "norm2_w.params",
        "post_attention_layernorm.bias": "norm2_b.params",
    }


class FalconNonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    lm_head: UnembedParameter
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "transformer.word_embeddings.weight": "embed.params",
        "transformer.ln_f.weight": "final_norm_w.params",
        "transformer.ln_f.bias": "final_norm_b.params",
        "lm_head.weight": "lm_head.params",
    }
