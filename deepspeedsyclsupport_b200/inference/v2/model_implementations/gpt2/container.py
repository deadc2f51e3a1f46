"""GPT-2: Conv1D modules store weights [in, out]; c_attn is q | k | v along the output axis; learned positions; tied unembedding."""
from ..common_parameters import (AttentionOutputParameter, EmbeddingParameter, FusedQKVParameter, MLP1Parameter, MLP2Parameter,
                                 NormParameter, UnembedParameter)
from ..layer_container_base import LayerContainer


class GPT2TransformerContainer(LayerContainer):
    qkv_w: FusedQKVParameter
    qkv_b: FusedQKVParameter
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
        "attn.c_attn.weight": ("qkv_w.params", "transpose"),
        "attn.c_attn.bias": "qkv_b.params",
        "attn.c_proj.weight": ("o_w.params", "transpose"),
        "attn.c_proj.bias": "o_b.params",
        "mlp.c_fc.weight": ("mlp1_w.params", "transpose"),
        "mlp.c_fc.bias": "mlp1_b.params",
        "mlp.c_proj.weight": ("mlp2_w.params", "transpose"),
        "mlp.c_proj.bias": "mlp2_b.params",
        "ln_1.weight": "norm1_w.params",
        "ln_1.bias": "norm1_b.params",
        "ln_2.weight": "norm2_w.params",
        "ln_2.bias": "norm2_b.params",
    }

    @staticmethod
    def transpose(t):
        return t.t().contiguous()


class GPT2NonTransformerContainer(LayerContainer):
    embed: EmbeddingParameter
    pos_embed: EmbeddingParameter
    lm_head: UnembedParameter
    final_norm_w: NormParameter
    final_norm_b: NormParameter

    PARAM_MAPPING = {
        "transformer.wte.weight": "embed.params",
        "transformer.wpe.weight": "pos_embed.params",
        "transformer.ln_f.weight": "final_norm_w.params",
        "transformer.ln_f.bias": "final_norm_b.params",
        "lm_head.weight": "lm_head.params",
    }
